// cvrng.h -- host-side generator of the Gaussian stream the reference draws from cv::RNG
// (cv::RNG rng(1), include/mantis3/Mantis3Params.h:87; rng.gaussian(sigma), PoseAdjustment.h:15-16).
// OpenCV's RNG is a multiply-with-carry generator (x <- lo(x) * 4164903690 + hi(x)) feeding a
// 128-strip Ziggurat for float32 normals; gaussian(sigma) widens the float and scales by sigma.
// The perturbation table of the refinement stage is built from it once per context.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdlib>

class CvRng {
public:
    explicit CvRng(uint64_t seed) : state_(seed ? seed : 0xffffffffULL) { init_tables(); }

    double gaussian(double sigma) { return (double)randn() * sigma; }

    float randn()
    {
        const float r = 3.442620f, rng_flt = 2.3283064365386962890625e-10f;  // 2^-32
        float x, y;
        for (;;) {
            int hz = (int)state_;
            next();
            int iz = hz & 127;
            x = hz * wn_[iz];
            if ((unsigned)std::abs(hz) < kn_[iz]) break;
            if (iz == 0) {  // tail
                do {
                    x = (unsigned)state_ * rng_flt; next();
                    y = (unsigned)state_ * rng_flt; next();
                    x = (float)(-std::log(x + FLT_MIN) * 0.2904764);
                    y = (float)-std::log(y + FLT_MIN);
                } while (y + y < x * x);
                x = hz > 0 ? r + x : -r - x;
                break;
            }
            y = (unsigned)state_ * rng_flt; next();
            if (fn_[iz] + y * (fn_[iz - 1] - fn_[iz]) < (float)std::exp(-.5 * x * x)) break;
        }
        return x;
    }

private:
    void next() { state_ = (uint64_t)(unsigned)state_ * 4164903690U + (state_ >> 32); }
    void init_tables()
    {
        const double m1 = 2147483648.0;
        double dn = 3.442619855899, tn = dn, vn = 9.91256303526217e-3;
        double q = vn / std::exp(-.5 * dn * dn);
        kn_[0] = (unsigned)((dn / q) * m1);
        kn_[1] = 0;
        wn_[0] = (float)(q / m1);
        wn_[127] = (float)(dn / m1);
        fn_[0] = 1.f;
        fn_[127] = (float)std::exp(-.5 * dn * dn);
        for (int i = 126; i >= 1; i--) {
            dn = std::sqrt(-2. * std::log(vn / dn + std::exp(-.5 * dn * dn)));
            kn_[i + 1] = (unsigned)((dn / tn) * m1);
            tn = dn;
            fn_[i] = (float)std::exp(-.5 * dn * dn);
            wn_[i] = (float)(dn / m1);
        }
    }
    uint64_t state_;
    unsigned kn_[128];
    float wn_[128], fn_[128];
};
