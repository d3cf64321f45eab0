// emul.cpp -- HOST build of the device RPP code (csrc/rpp_fast.cuh + rpp.cuh), a diagnostic aid:
// the orthogonal iteration is a chain of dependent fp64 steps whose STOP decision
// (RPP.cpp:171, |(old-new)/old| > 1e-5) must fall on the same iteration as the oracle's, and that can
// be studied on the CPU, quad by quad, without a GPU.  Mirrors k_pose_run / k_pose_second /
// k_pose_select of csrc/pose.cu.  Never linked into libmantis_cuda.so.
//   g++ -O2 -std=c++17 -ffp-contract=off -mfma -shared -fPIC -o librpp_emul.so emul.cpp
#include <math.h>
#include <string.h>
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
#define MANTIS_HOST_EMUL 1
#include "../../mantis_b200/csrc/rpp_fast.cuh"

namespace {
void model_xy(double s, int ordering, double* m)
{
    const double sx[2][4] = {{1, -1, -1, 1}, {1, -1, -1, 1}}, sy[2][4] = {{1, 1, -1, -1}, {-1, -1, 1, 1}};
    for (int k = 0; k < 4; ++k) { m[2 * k] = sx[ordering][k] * s; m[2 * k + 1] = sy[ordering][k] * s; }
}
struct RunOut { double R[9], t[3], obj_err, img_err; int it; };

RunOut run(const double* rays, const double* m, const double* initR)
{
    const double TOL = 1e-5, EPSILON = 1e-8;
    double ctx[rppf::CTX_WORDS];
    rppf::Lane L;
    rppf::setup(ctx, 1, rays, m, L);
    rppf::Run S;
    rppf::run_begin(ctx, 1, L, S, initR);
    while (!rppf::run_step(ctx, 1, L, S, TOL, EPSILON)) {}
    RunOut o;
    rppf::finalize(ctx, 1, L, o.R, o.t, o.obj_err, o.img_err);
    o.it = S.it;
    return o;
}
}  // namespace

// rays8: 4 undistorted test points; returns 1 on success.  iters[0] = first run, iters[1..] = candidate runs.
extern "C" int emul_rpp(const double* rays8, double half_spacing, int ordering, double* R9, double* t3, int* iters, int* n_iters,
                        double* obj_err, double* img_err)
{
    double m[8];
    model_xy(half_spacing, ordering, m);
    RunOut r1 = run(rays8, m, nullptr);
    iters[0] = r1.it;
    *n_iters = 1;
    pm::V3 model[4], ray[4];
    for (int k = 0; k < 4; ++k) {
        model[k].v[0] = m[2 * k]; model[k].v[1] = m[2 * k + 1]; model[k].v[2] = 0;
        ray[k].v[0] = rays8[2 * k]; ray[k].v[1] = rays8[2 * k + 1]; ray[k].v[2] = 1.0;
    }
    pm::M3 R, solR[4];
    pm::V3 solT[4];
    memcpy(R.m, r1.R, sizeof(R.m));
    const int ns = rpp::second_poses<4>(model, ray, R, solR, solT);
    if (ns <= 0) return 0;
    int best = -1;
    double lowest = 1e6;
    RunOut bo{};
    for (int c = 0; c < ns; ++c) {
        RunOut r2 = run(rays8, m, solR[c].m);
        iters[(*n_iters)++] = r2.it;
        if (r2.obj_err < lowest) { lowest = r2.obj_err; best = c; bo = r2; }
    }
    if (best < 0) return 0;
    memcpy(R9, bo.R, sizeof(bo.R));
    memcpy(t3, bo.t, sizeof(bo.t));
    *obj_err = bo.obj_err;
    *img_err = bo.img_err;
    return 1;
}
