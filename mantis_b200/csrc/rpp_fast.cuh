// rpp_fast.cuh -- the object-space orthogonal iteration of RPP (ObjPose / AbsKernel,
// include/mantis3/RobustPlanarPose/RPP.cpp:66-332) for a PLANAR 4-point model, one iteration at a
// time, so that a warp can interleave many independent runs (pose.cu's work-stealing loop).
//
// Planarity makes the absolute-orientation step closed form: with centred model points
// P_i = (px_i, py_i, 0), M = sum P_i Q_i^T has a zero third row, so the rotation V U^T of its SVD
// is [r1 r2 r1 x r2] with [r1 r2] the orthonormal polar factor of A = [a b] = M^T[:, 0:2]:
//     [r1 r2] = A (A^T A)^-1/2,   (A^T A)^-1/2 = [[S22 + s, -S12], [-S12, S11 + s]] / (k s),
//     s = |a x b| = sqrt(det A^T A),  k = sqrt(trace + 2 s).
// The reference's sign fix-ups (RPP.cpp:283-310) only act when the translation comes out behind
// the camera; that case, and a numerically rank-deficient A, go through the general SVD step
// (rpp::abs_kernel in rpp.cuh).
//
// Between two steps the points are Q_i = R P_i + t, so everything a step needs is linear or
// quadratic in (r1, r2, t) with per-run constant matrices (F_i = v_i v_i^T / v_i.v_i, symmetric):
//     a = sum px_i F_i Q_i = Gxx r1 + Gxy r2 + Gx t        Gxx = sum px_i^2 F_i, Gxy = sum px_i py_i F_i,
//     b = sum py_i F_i Q_i = Gxy r1 + Gyy r2 + Gy t        Gx = sum px_i F_i, G1 = sum F_i, ...
//     t = tF sum F_i R P_i = tF (Gx r1 + Gy r2)
//     sum |Q_i - F_i Q_i|^2 = sum |Q_i|^2 - (r1.a + r2.b + t.(Gx r1 + Gy r2 + G1 t))     (I - F_i is a projector)
// and, t being linear in (r1, r2) and (r1, r2) linear in (a, b) once the three polar coefficients c
// are known,
//     a' = c11 Kxx a + c12 (Kxx b + Kxy a) + c22 Kxy b      Kxx = Gxx + Gx tF Gx,  Kxy = Gxy + Gx tF Gy,
//     b' = c11 Kxy^T a + c12 (Kxy^T b + Kyy a) + c22 Kyy b  Kyy = Gyy + Gy tF Gy.
// A run is a chain of dependent steps and the longest runs (thousands of steps) set the duration of
// the whole launch.  The a', b' form above lets every matrix-vector product issue beside the scalar chain
// (a, b) -> c, but costs twelve products per step; the step is bound by the fp64 pipe, not by that chain, so
// advance_fast() forms (r1, r2) first and multiplies five matrices.  The state of a run is (a, b, r1, r2, t).
//
// The error that feeds the stopping rule (RPP.cpp:171, |(old - new) / old| > 1e-5) is accumulated point
// by point from (R, t), as the reference does.  The closed quadratic form above would be cheaper, but it
// carries the cancellation of that difference (~1e-14 absolute against errors down to 1e-8, i.e. 1e-6
// relative, a tenth of the tolerance): measured on the quads of 1024 rendered frames
// (tools/rpp_emul/compare.py) it moved the stop by one or more iterations in 30 of 112 790 solves, in
// 15 of the 1024 frames -- and a run that stops one step apart hands the scoring stage a pose 1e-6
// away, enough to flip a pixel rounding and send the greedy refinement down another path.  With the
// point-by-point error the same experiment gives 0 of 112 790.  The extra work (4 points x ~16 fp64
// operations) is off the dependent chain.
#pragma once
#include "fastmath.cuh"
#include "rpp.cuh"

namespace rppf {

constexpr int NPT = 4;

// Read-only data of one run, kept in shared memory column `lane` of a [CTX_WORDS][threads] array.
// Symmetric matrices are stored as xx xy xz yy yz zz.
// Kxy is general (row-major, 9 words).
enum {
    C_VX = 0, C_VY = 4, C_PX = 8, C_PY = 12, C_KXX = 16, C_KXY = 22, C_KYY = 31, C_GX = 37, C_GY = 43, C_IVV = 49, C_TF = 53,
    C_PBX = 59, C_PBY = 60, CTX_WORDS = 61
};

struct Lane {            // registers of one run
    double a[3], b[3];   // sum px_i F_i Q_i, sum py_i F_i Q_i for the current points
    double r1[3], r2[3], t[3];
    bool rays;           // the current points are still the image rays (first step of a run without initial R)
};

#define RPPF_C(k) ctx[(k) * stride]

// out = G x for the symmetric matrix at context word g
__device__ __forceinline__ void symv(const double* ctx, int stride, int g, const double* x, double* out)
{
    const double xx = RPPF_C(g), xy = RPPF_C(g + 1), xz = RPPF_C(g + 2), yy = RPPF_C(g + 3), yz = RPPF_C(g + 4), zz = RPPF_C(g + 5);
    out[0] = fma(xz, x[2], fma(xy, x[1], xx * x[0]));
    out[1] = fma(yz, x[2], fma(yy, x[1], xy * x[0]));
    out[2] = fma(zz, x[2], fma(yz, x[1], xz * x[0]));
}
__device__ __forceinline__ void symv_acc(const double* ctx, int stride, int g, const double* x, double* out)
{
    const double xx = RPPF_C(g), xy = RPPF_C(g + 1), xz = RPPF_C(g + 2), yy = RPPF_C(g + 3), yz = RPPF_C(g + 4), zz = RPPF_C(g + 5);
    out[0] = fma(xz, x[2], fma(xy, x[1], fma(xx, x[0], out[0])));
    out[1] = fma(yz, x[2], fma(yy, x[1], fma(xy, x[0], out[1])));
    out[2] = fma(zz, x[2], fma(yz, x[1], fma(xz, x[0], out[2])));
}

// out = K x and out_t = K^T y for the general matrix at context word g
__device__ __forceinline__ void genv2(const double* ctx, int stride, int g, const double* x, const double* y, double* out, double* out_t)
{
    double k[9];
    for (int i = 0; i < 9; ++i) k[i] = RPPF_C(g + i);
    for (int r = 0; r < 3; ++r) {
        out[r] = fma(k[3 * r + 2], x[2], fma(k[3 * r + 1], x[1], k[3 * r] * x[0]));
        out_t[r] = fma(k[6 + r], y[2], fma(k[3 + r], y[1], k[r] * y[0]));
    }
}

// rays (x, y, 1) and model (x, y, 0) -> context; the run starts from Q_i = v_i.  Mirrors rpp::setup.
__device__ inline void setup(double* ctx, int stride, const double* ray_xy, const double* model_xy, Lane& L)
{
    double pbx = 0, pby = 0;
    for (int i = 0; i < NPT; ++i) { pbx += model_xy[2 * i]; pby += model_xy[2 * i + 1]; }
    pbx /= NPT; pby /= NPT;
    double gxx[6], gxy[6], gyy[6], gx[6], gy[6], g1[6];
    for (int k = 0; k < 6; ++k) gxx[k] = gxy[k] = gyy[k] = gx[k] = gy[k] = g1[k] = 0;
    for (int k = 0; k < 3; ++k) L.a[k] = L.b[k] = L.r1[k] = L.r2[k] = L.t[k] = 0;
    for (int i = 0; i < NPT; ++i) {
        const double v[3] = {ray_xy[2 * i], ray_xy[2 * i + 1], 1.0};
        const double vv = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
        const double px = model_xy[2 * i] - pbx, py = model_xy[2 * i + 1] - pby;
        RPPF_C(C_VX + i) = v[0];
        RPPF_C(C_VY + i) = v[1];
        RPPF_C(C_PX + i) = px;
        RPPF_C(C_PY + i) = py;
        RPPF_C(C_IVV + i) = 1.0 / vv;
        const double f[6] = {(v[0] * v[0]) / vv, (v[0] * v[1]) / vv, (v[0] * v[2]) / vv, (v[1] * v[1]) / vv, (v[1] * v[2]) / vv, (v[2] * v[2]) / vv};
        for (int k = 0; k < 6; ++k) {
            g1[k] += f[k];
            gx[k] += px * f[k];
            gy[k] += py * f[k];
            gxx[k] += px * px * f[k];
            gxy[k] += px * py * f[k];
            gyy[k] += py * py * f[k];
        }
        for (int k = 0; k < 3; ++k) { L.a[k] += px * v[k]; L.b[k] += py * v[k]; }  // F_i v_i = v_i
    }
    L.rays = true;
    pm::M3 A = pm::eye();
    const int full[9] = {0, 1, 2, 1, 3, 4, 2, 4, 5};
    for (int k = 0; k < 9; ++k) A.m[k] -= g1[full[k]] / NPT;
    pm::M3 tF = pm::inverse(A);
    const int upper[6] = {0, 1, 2, 4, 5, 8};
    double tf[9], GX[9], GY[9], TX[9], TY[9];   // tF, Gx, Gy, tF Gx, tF Gy as full matrices
    for (int k = 0; k < 9; ++k) { tf[k] = tF.m[upper[full[k]]] / NPT; GX[k] = gx[full[k]]; GY[k] = gy[full[k]]; }
    for (int r = 0; r < 3; ++r)
        for (int cc = 0; cc < 3; ++cc) {
            TX[3 * r + cc] = tf[3 * r] * GX[cc] + tf[3 * r + 1] * GX[3 + cc] + tf[3 * r + 2] * GX[6 + cc];
            TY[3 * r + cc] = tf[3 * r] * GY[cc] + tf[3 * r + 1] * GY[3 + cc] + tf[3 * r + 2] * GY[6 + cc];
        }
    for (int r = 0; r < 3; ++r)
        for (int cc = 0; cc < 3; ++cc) {
            const double xx = GX[3 * r] * TX[cc] + GX[3 * r + 1] * TX[3 + cc] + GX[3 * r + 2] * TX[6 + cc];
            const double xy = GX[3 * r] * TY[cc] + GX[3 * r + 1] * TY[3 + cc] + GX[3 * r + 2] * TY[6 + cc];
            const double yy = GY[3 * r] * TY[cc] + GY[3 * r + 1] * TY[3 + cc] + GY[3 * r + 2] * TY[6 + cc];
            RPPF_C(C_KXY + 3 * r + cc) = gxy[full[3 * r + cc]] + xy;
            if (cc >= r) {
                const int u = r == 0 ? cc : (r == 1 ? 2 + cc : 5);
                RPPF_C(C_KXX + u) = gxx[u] + xx;
                RPPF_C(C_KYY + u) = gyy[u] + yy;
            }
        }
    for (int k = 0; k < 6; ++k) {
        RPPF_C(C_TF + k) = tF.m[upper[k]] / NPT;
        RPPF_C(C_GX + k) = gx[k]; RPPF_C(C_GY + k) = gy[k];
    }
    RPPF_C(C_PBX) = pbx;
    RPPF_C(C_PBY) = pby;
}

// The current points of the run: the rays, or R P_i + t.
__device__ __forceinline__ void point(const double* ctx, int stride, const Lane& L, int i, double* q)
{
    if (L.rays) { q[0] = RPPF_C(C_VX + i); q[1] = RPPF_C(C_VY + i); q[2] = 1.0; return; }
    const double px = RPPF_C(C_PX + i), py = RPPF_C(C_PY + i);
    for (int k = 0; k < 3; ++k) q[k] = (L.r1[k] * px + L.r2[k] * py) + L.t[k];
}

// sum |Q_i - F_i Q_i|^2 point by point for Q_i = R P_i + t, as the reference accumulates it
// (F_i Q_i = v_i (v_i.Q_i) / (v_i.v_i))
__device__ __forceinline__ double point_error_rt(const double* ctx, int stride, const double* r1, const double* r2, const double* t)
{
    double err2 = 0;
    for (int i = 0; i < NPT; ++i) {
        const double px = RPPF_C(C_PX + i), py = RPPF_C(C_PY + i), vx = RPPF_C(C_VX + i), vy = RPPF_C(C_VY + i);
        const double q0 = fma(r2[0], py, fma(r1[0], px, t[0])), q1 = fma(r2[1], py, fma(r1[1], px, t[1])),
                     q2 = fma(r2[2], py, fma(r1[2], px, t[2]));
        const double w = fma(vy, q1, fma(vx, q0, q2)) * RPPF_C(C_IVV + i);
        const double e0 = fma(-vx, w, q0), e1 = fma(-vy, w, q1), e2 = q2 - w;
        err2 = fma(e2, e2, fma(e1, e1, fma(e0, e0, err2)));
    }
    return err2;
}
__device__ inline double point_error(const double* ctx, int stride, const Lane& L)
{
    if (!L.rays) return point_error_rt(ctx, stride, L.r1, L.r2, L.t);
    return 0;  // the rays lie on their own lines of sight
}

// With (r1, r2) set and t = tF (Gx r1 + Gy r2): t, a, b of the new points and their error.
__device__ inline double refresh(const double* ctx, int stride, Lane& L)
{
    double gs[3], ta[3], tb[3];
    symv(ctx, stride, C_GX, L.r1, gs);
    symv_acc(ctx, stride, C_GY, L.r2, gs);
    symv(ctx, stride, C_TF, gs, L.t);
    symv(ctx, stride, C_KXX, L.r1, L.a);
    symv(ctx, stride, C_KYY, L.r2, L.b);
    genv2(ctx, stride, C_KXY, L.r2, L.r1, ta, tb);
    for (int k = 0; k < 3; ++k) { L.a[k] += ta[k]; L.b[k] += tb[k]; }
    L.rays = false;
    return point_error_rt(ctx, stride, L.r1, L.r2, L.t);
}

// The general step through the 3x3 SVD (rare): rebuilds the reference-shaped context.
__device__ __noinline__ double abs_kernel_general_impl(const double* ctx, int stride, Lane& L);
// (through a copy, so that the caller's Lane never has its address taken and stays in registers)
__device__ __forceinline__ double abs_kernel_general(const double* ctx, int stride, Lane& L)
{
    Lane tmp = L;
    const double e = abs_kernel_general_impl(ctx, stride, tmp);
    L = tmp;
    return e;
}
__device__ __noinline__ double abs_kernel_general_impl(const double* ctx, int stride, Lane& L)
{
    rpp::ObjCtx<NPT> c;
    pm::V3 Q[NPT];
    for (int i = 0; i < NPT; ++i) {
        const double v[3] = {RPPF_C(C_VX + i), RPPF_C(C_VY + i), 1.0};
        const double vv = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
        c.P[i].v[0] = RPPF_C(C_PX + i); c.P[i].v[1] = RPPF_C(C_PY + i); c.P[i].v[2] = 0;
        for (int a = 0; a < 3; ++a) {
            c.Qp[i].v[a] = v[a];
            for (int b = 0; b < 3; ++b) c.F[i].m[a * 3 + b] = (v[a] * v[b]) / vv;
        }
        point(ctx, stride, L, i, Q[i].v);
    }
    const int full[9] = {0, 1, 2, 1, 3, 4, 2, 4, 5};
    for (int k = 0; k < 9; ++k) c.tF.m[k] = RPPF_C(C_TF + full[k]);
    c.pbar.v[0] = RPPF_C(C_PBX); c.pbar.v[1] = RPPF_C(C_PBY); c.pbar.v[2] = 0;
    pm::M3 R;
    pm::V3 t;
    double err2;
    rpp::abs_kernel<NPT>(c, Q, R, t, err2);
    for (int k = 0; k < 3; ++k) { L.r1[k] = R.m[3 * k]; L.r2[k] = R.m[3 * k + 1]; }
    refresh(ctx, stride, L);  // (t = tF sum F_i R P_i there as in the reference step)
    return err2;
}

// The closed-form part of one AbsKernel step (RPP.cpp:229-332): from the current points (through L.a, L.b) to
// the next R = [r1 r2 r1 x r2], t and the (a, b) of the points they give.  Nothing of L is changed; false = this step
// needs the general SVD step (rank-deficient A, or a translation behind the camera).
// After the polar coefficients c the step is refresh() on the new (r1, r2): five matrix-vector products (t, a', b')
// instead of the twelve that the a / b-only formulation issues beside the scalar chain -- that formulation is ~50
// cycles shorter in latency, but the instruction count is what a step costs.
__device__ __forceinline__ bool advance_fast(const double* ctx, int stride, const Lane& L, double* r1, double* r2, double* t,
                                             double* an, double* bn)
{
    const double* a = L.a;
    const double* b = L.b;
    const double S11 = fma(a[2], a[2], fma(a[1], a[1], a[0] * a[0])), S22 = fma(b[2], b[2], fma(b[1], b[1], b[0] * b[0])),
                 S12 = fma(a[2], b[2], fma(a[1], b[1], a[0] * b[0]));
    const double n0 = fma(a[1], b[2], -(a[2] * b[1])), n1 = fma(a[2], b[0], -(a[0] * b[2])), n2 = fma(a[0], b[1], -(a[1] * b[0]));
    const double nn = fma(n2, n2, fma(n1, n1, n0 * n0)), tr = S11 + S22;
    const bool full_rank = nn > 1e-20 * (tr * tr);   // else s <= 1e-10 tr
    // 1/s and 1/k from the reciprocal-square-root seed + two Newton steps (no sqrt, no division)
    const double inv_s = fast_rsqrt(nn), s = nn * inv_s;
    const double inv = fast_rsqrt(fma(2.0, s, tr)) * inv_s;
    const double c11 = (S22 + s) * inv, c12 = -S12 * inv, c22 = (S11 + s) * inv;
    double gs[3], ta[3], tb[3];
    for (int k = 0; k < 3; ++k) {
        r1[k] = fma(b[k], c12, a[k] * c11);
        r2[k] = fma(b[k], c22, a[k] * c12);
    }
    symv(ctx, stride, C_GX, r1, gs);
    symv_acc(ctx, stride, C_GY, r2, gs);
    symv(ctx, stride, C_TF, gs, t);   // t = tF * sum F_i (R P_i)
    symv(ctx, stride, C_KXX, r1, an);
    symv(ctx, stride, C_KYY, r2, bn);
    genv2(ctx, stride, C_KXY, r2, r1, ta, tb);
    for (int k = 0; k < 3; ++k) { an[k] += ta[k]; bn[k] += tb[k]; }
    return full_rank && !(t[2] < 0);
}

// ObjPose with an initial rotation (RPP.cpp:98-128): t from R, Q_i, and the starting error.
// (sum (F_i - I) R P_i = Gx r1 + Gy r2 because the model points are centred.)
__device__ inline double init_from_R(const double* ctx, int stride, const double* R9, Lane& L)
{
    for (int k = 0; k < 3; ++k) { L.r1[k] = R9[3 * k]; L.r2[k] = R9[3 * k + 1]; }
    return refresh(ctx, stride, L);
}

// End of ObjPose (RPP.cpp:176-207): R, t (un-centred), object- and image-space errors.
__device__ inline void finalize(const double* ctx, int stride, const Lane& L, double* R9, double* t3, double& obj_err, double& img_err)
{
    const double r3[3] = {L.r1[1] * L.r2[2] - L.r1[2] * L.r2[1], L.r1[2] * L.r2[0] - L.r1[0] * L.r2[2], L.r1[0] * L.r2[1] - L.r1[1] * L.r2[0]};
    for (int k = 0; k < 3; ++k) { R9[3 * k] = L.r1[k]; R9[3 * k + 1] = L.r2[k]; R9[3 * k + 2] = r3[k]; }
    obj_err = sqrt(point_error(ctx, stride, L) / NPT);
    double ie = 0;
    const double u0 = RPPF_C(C_VX + 0), v0 = RPPF_C(C_VY + 0);
    for (int i = 0; i < NPT; ++i) {
        double q[3];
        point(ctx, stride, L, i, q);
        // reference quirk (RPP.cpp:197-198): every point is compared with image point 0
        const double xx = q[0] / q[2] - u0, yy = q[1] / q[2] - v0;
        ie += xx * xx + yy * yy;
    }
    img_err = sqrt(ie / NPT);
    const double pbx = RPPF_C(C_PBX), pby = RPPF_C(C_PBY);
    for (int k = 0; k < 3; ++k) t3[k] = L.t[k] - (L.r1[k] * pbx + L.r2[k] * pby);
}

// ---- one ObjPose run as a resumable state machine (RPP.cpp:66-208) --------------------------------
// run_begin() after setup(); then run_step() until it returns true; then finalize().
//
// The reference's loop is  step k -> error of step k -> stop test -> step k + 1,  one dependent chain of ~60 fp64
// operations per step, and a lone warp in the tail of the launch issues one of them every ~8 cycles.  Here a trip
// of run_step() computes the error of the COMMITTED step k - 1 and, beside it, step k from the same state (the
// two do not depend on each other); the stop test of step k - 1 then decides whether step k is committed or
// thrown away.  The tests, their order and the state a run ends in are the reference's; a run costs one trip more.
struct Run {
    double old_err, new_err;
    double e_known;  // error of the committed step when it came with the step (general SVD step)
    int it;          // stop tests made so far
    int pending;     // 2: the next error is old_err (run without initial R), 1: the next error is the first new_err
    bool have;       // L.(r1, r2, t) is a step whose error has not been looked at yet
    bool known;      // ... and that error is e_known
};
__device__ __forceinline__ void run_begin(const double* ctx, int stride, Lane& L, Run& S, const double* initR9)
{
    S.it = 0;
    S.new_err = 0;
    S.e_known = 0;
    S.have = false;
    S.known = false;
    if (initR9) {
        S.old_err = init_from_R(ctx, stride, initR9, L);
        S.pending = 1;
    } else {
        S.old_err = 0;
        S.pending = 2;  // the run starts from the rays themselves (setup)
    }
}
// One trip: the reference's stopping rule on the committed step, and the next AbsKernel step.  Returns true when
// the run is over (L then holds the step the rule stopped at).
__device__ __forceinline__ bool run_step(const double* ctx, int stride, Lane& L, Run& S, double TOL, double EPSILON)
{
    const double e_calc = point_error_rt(ctx, stride, L.r1, L.r2, L.t);   // (of no use while !S.have: finite garbage)
    double r1[3], r2[3], t[3], an[3], bn[3];
    const bool ok = advance_fast(ctx, stride, L, r1, r2, t, an, bn);
    if (S.have) {
        const double e = S.known ? S.e_known : e_calc;
        if (S.pending == 2) {
            S.old_err = e;
            S.pending = 1;
        } else {
            if (S.pending == 0) S.old_err = S.new_err;
            S.new_err = e;
            S.pending = 0;
            // |(old - new) / old| > TOL without the division
            if (!(fabs(S.old_err - S.new_err) > TOL * fabs(S.old_err) && S.new_err > EPSILON) || ++S.it > 100000) return true;
        }
    }
    if (ok) {
        for (int k = 0; k < 3; ++k) { L.r1[k] = r1[k]; L.r2[k] = r2[k]; L.t[k] = t[k]; L.a[k] = an[k]; L.b[k] = bn[k]; }
        L.rays = false;
        S.known = false;
    } else {
        S.e_known = abs_kernel_general(ctx, stride, L);
        S.known = true;
    }
    S.have = true;
    return false;
}

#undef RPPF_C
}  // namespace rppf
