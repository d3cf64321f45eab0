/*
 * camera_rng.c -- oracle (test infrastructure) for the OpenCV routines the reference calls
 * around its geometry:
 *   cv::fisheye::undistortPoints(pts, pts, K, D)   include/mantis3/QuadDetection.h:294
 *   cv::fisheye::distortPoints(undist, dist, K, D) include/mantis3/Mantis3Types.h:129
 *   cv::RNG::gaussian(sigma)                       include/mantis3/PoseAdjustment.h:15-16
 *                                                  (cv::RNG rng(1), Mantis3Params.h:87)
 * OpenCV is not vendored by the reference; these restate OpenCV 4.13's published
 * algorithms (Kannala-Brandt theta polynomial with Newton inversion; MWC generator with a
 * 128-strip Ziggurat) and are pinned against cv2 4.13.0 by tests/test_oracle_camera_rng.py.
 */
#include "mantis_oracle.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>

/* K: 3x3 row-major FLOAT (the reference builds a CV_32F K: get3x3FromVector,
 * QuadDetection.h:189-201); D: 4 doubles. */
void mo_fisheye_undistort(const double* src_xy, int n, const float* K, const double* D, double* dst_xy)
{
    const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    const double eps = 1e-8;
    for (int i = 0; i < n; ++i) {
        double pwx = (src_xy[2 * i] - cx) / fx, pwy = (src_xy[2 * i + 1] - cy) / fy;
        double theta_d = sqrt(pwx * pwx + pwy * pwy);
        theta_d = fmin(fmax(-M_PI / 2., theta_d), M_PI / 2.);
        int converged = 0;
        double theta = theta_d, scale = 0.0;
        if (fabs(theta_d) > eps) {
            for (int j = 0; j < 10; ++j) {
                double t2 = theta * theta, t4 = t2 * t2, t6 = t4 * t2, t8 = t6 * t2;
                double k0t2 = D[0] * t2, k1t4 = D[1] * t4, k2t6 = D[2] * t6, k3t8 = D[3] * t8;
                double fix = (theta * (1 + k0t2 + k1t4 + k2t6 + k3t8) - theta_d) /
                             (1 + 3 * k0t2 + 5 * k1t4 + 7 * k2t6 + 9 * k3t8);
                theta = theta - fix;
                if (fabs(fix) < eps) { converged = 1; break; }
            }
            scale = tan(theta) / theta_d;
        } else {
            converged = 1;
        }
        int flipped = (theta_d < 0 && theta > 0) || (theta_d > 0 && theta < 0);
        if (converged && !flipped) {
            dst_xy[2 * i] = pwx * scale;
            dst_xy[2 * i + 1] = pwy * scale;
        } else {
            dst_xy[2 * i] = -1000000.0;
            dst_xy[2 * i + 1] = -1000000.0;
        }
    }
}

void mo_fisheye_distort(const double* src_xy, int n, const float* K, const double* D, double* dst_xy)
{
    const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    for (int i = 0; i < n; ++i) {
        double x = src_xy[2 * i], y = src_xy[2 * i + 1];
        double r2 = x * x + y * y;
        double r = sqrt(r2);
        double theta = atan(r);
        double t2 = theta * theta, t3 = t2 * theta, t4 = t2 * t2, t5 = t4 * theta, t6 = t3 * t3, t7 = t6 * theta,
               t8 = t4 * t4, t9 = t8 * theta;
        double theta_d = theta + D[0] * t3 + D[1] * t5 + D[2] * t7 + D[3] * t9;
        double inv_r = r > 1e-8 ? 1.0 / r : 1;
        double cdist = r > 1e-8 ? theta_d * inv_r : 1;
        double xd = x * cdist, yd = y * cdist;
        dst_xy[2 * i] = xd * fx + cx;
        dst_xy[2 * i + 1] = yd * fy + cy;
    }
}

/* ---- cv::RNG: multiply-with-carry + Ziggurat normal ---- */
static unsigned zig_kn[128];
static float zig_wn[128], zig_fn[128];
static int zig_init = 0;

static void zig_setup(void)
{
    const double m1 = 2147483648.0;
    double dn = 3.442619855899, tn = dn, vn = 9.91256303526217e-3;
    double q = vn / exp(-.5 * dn * dn);
    zig_kn[0] = (unsigned)((dn / q) * m1);
    zig_kn[1] = 0;
    zig_wn[0] = (float)(q / m1);
    zig_wn[127] = (float)(dn / m1);
    zig_fn[0] = 1.f;
    zig_fn[127] = (float)exp(-.5 * dn * dn);
    for (int i = 126; i >= 1; i--) {
        dn = sqrt(-2. * log(vn / dn + exp(-.5 * dn * dn)));
        zig_kn[i + 1] = (unsigned)((dn / tn) * m1);
        tn = dn;
        zig_fn[i] = (float)exp(-.5 * dn * dn);
        zig_wn[i] = (float)(dn / m1);
    }
    zig_init = 1;
}

void mo_ziggurat_tables(unsigned* kn, float* wn, float* fn)
{
    if (!zig_init) zig_setup();
    for (int i = 0; i < 128; ++i) { kn[i] = zig_kn[i]; wn[i] = zig_wn[i]; fn[i] = zig_fn[i]; }
}

#define RNG_NEXT(x) ((uint64_t)(unsigned)(x) * 4164903690U + ((x) >> 32))

/* one standard-normal float32 draw, advancing *state (cv::RNG state; seed 0 -> 0xffffffff) */
float mo_rng_randn(uint64_t* state)
{
    if (!zig_init) zig_setup();
    const float r = 3.442620f;
    const float rng_flt = 2.3283064365386962890625e-10f;
    uint64_t temp = *state;
    float x, y;
    for (;;) {
        int hz = (int)temp;
        temp = RNG_NEXT(temp);
        int iz = hz & 127;
        x = hz * zig_wn[iz];
        if ((unsigned)abs(hz) < zig_kn[iz]) break;
        if (iz == 0) {
            do {
                x = (unsigned)temp * rng_flt;
                temp = RNG_NEXT(temp);
                y = (unsigned)temp * rng_flt;
                temp = RNG_NEXT(temp);
                x = (float)(-log(x + FLT_MIN) * 0.2904764);
                y = (float)-log(y + FLT_MIN);
            } while (y + y < x * x);
            x = hz > 0 ? r + x : -r - x;
            break;
        }
        y = (unsigned)temp * rng_flt;
        temp = RNG_NEXT(temp);
        if (zig_fn[iz] + y * (zig_fn[iz - 1] - zig_fn[iz]) < (float)exp(-.5 * x * x)) break;
    }
    *state = temp;
    return x;
}

/* cv::RNG::gaussian(sigma): float32 normal widened and scaled in double */
double mo_rng_gaussian(uint64_t* state, double sigma) { return (double)mo_rng_randn(state) * sigma; }
