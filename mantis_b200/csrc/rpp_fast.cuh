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
// (rpp::abs_kernel in rpp.cuh).  F_i x is evaluated as v_i (v_i . x) / (v_i . v_i).
#pragma once
#include "rpp.cuh"

namespace rppf {

constexpr int NPT = 4;

// Read-only data of one run, kept in shared memory column `lane` of a [CTX_WORDS][threads] array.
enum { C_VX = 0, C_VY = 4, C_IVV = 8, C_PX = 12, C_PY = 16, C_TF = 20, C_PBX = 29, C_PBY = 30, CTX_WORDS = 31 };

struct Lane {            // registers of one run
    double Q[NPT][3];
    double r1[3], r2[3], r3[3], t[3];
};

#define RPPF_C(k) ctx[(k) * stride]

// rays (x, y, 1) and model (x, y, 0) -> context.  Mirrors rpp::setup.
__device__ inline void setup(double* ctx, int stride, const double* ray_xy, const double* model_xy)
{
    double pbx = 0, pby = 0;
    for (int i = 0; i < NPT; ++i) { pbx += model_xy[2 * i]; pby += model_xy[2 * i + 1]; }
    pbx /= NPT; pby /= NPT;
    pm::M3 sumF;
    for (int k = 0; k < 9; ++k) sumF.m[k] = 0;
    for (int i = 0; i < NPT; ++i) {
        const double v[3] = {ray_xy[2 * i], ray_xy[2 * i + 1], 1.0};
        const double vv = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
        RPPF_C(C_VX + i) = v[0];
        RPPF_C(C_VY + i) = v[1];
        RPPF_C(C_IVV + i) = 1.0 / vv;
        RPPF_C(C_PX + i) = model_xy[2 * i] - pbx;
        RPPF_C(C_PY + i) = model_xy[2 * i + 1] - pby;
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) sumF.m[a * 3 + b] += (v[a] * v[b]) / vv;
    }
    pm::M3 A = pm::eye();
    for (int k = 0; k < 9; ++k) A.m[k] -= sumF.m[k] / NPT;
    pm::M3 tF = pm::inverse(A);
    for (int k = 0; k < 9; ++k) RPPF_C(C_TF + k) = tF.m[k] / NPT;
    RPPF_C(C_PBX) = pbx;
    RPPF_C(C_PBY) = pby;
}

// F_i x
__device__ __forceinline__ void apply_F(const double* ctx, int stride, int i, const double* x, double* out)
{
    const double vx = RPPF_C(C_VX + i), vy = RPPF_C(C_VY + i);
    const double w = (vx * x[0] + vy * x[1] + x[2]) * RPPF_C(C_IVV + i);
    out[0] = vx * w; out[1] = vy * w; out[2] = w;
}

__device__ __forceinline__ void mul_tF(const double* ctx, int stride, const double* s, double* t)
{
    for (int r = 0; r < 3; ++r)
        t[r] = RPPF_C(C_TF + 3 * r) * s[0] + RPPF_C(C_TF + 3 * r + 1) * s[1] + RPPF_C(C_TF + 3 * r + 2) * s[2];
}

// Q_i = R P_i + t; returns sum |Q_i - F_i Q_i|^2
__device__ __forceinline__ double reproject(const double* ctx, int stride, Lane& L)
{
    double err2 = 0;
    for (int i = 0; i < NPT; ++i) {
        const double px = RPPF_C(C_PX + i), py = RPPF_C(C_PY + i);
        double q[3], fq[3];
        for (int k = 0; k < 3; ++k) q[k] = (L.r1[k] * px + L.r2[k] * py) + L.t[k];
        apply_F(ctx, stride, i, q, fq);
        for (int k = 0; k < 3; ++k) { L.Q[i][k] = q[k]; double e = q[k] - fq[k]; err2 += e * e; }
    }
    return err2;
}

// The general step through the 3x3 SVD (rare): rebuilds the reference-shaped context.
__device__ __noinline__ double abs_kernel_general(const double* ctx, int stride, Lane& L)
{
    rpp::ObjCtx<NPT> c;
    for (int i = 0; i < NPT; ++i) {
        const double v[3] = {RPPF_C(C_VX + i), RPPF_C(C_VY + i), 1.0};
        const double vv = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
        c.P[i].v[0] = RPPF_C(C_PX + i); c.P[i].v[1] = RPPF_C(C_PY + i); c.P[i].v[2] = 0;
        for (int a = 0; a < 3; ++a) {
            c.Qp[i].v[a] = v[a];
            for (int b = 0; b < 3; ++b) c.F[i].m[a * 3 + b] = (v[a] * v[b]) / vv;
        }
    }
    for (int k = 0; k < 9; ++k) c.tF.m[k] = RPPF_C(C_TF + k);
    c.pbar.v[0] = RPPF_C(C_PBX); c.pbar.v[1] = RPPF_C(C_PBY); c.pbar.v[2] = 0;
    pm::V3 Q[NPT];
    for (int i = 0; i < NPT; ++i)
        for (int k = 0; k < 3; ++k) Q[i].v[k] = L.Q[i][k];
    pm::M3 R;
    pm::V3 t;
    double err2;
    rpp::abs_kernel<NPT>(c, Q, R, t, err2);
    for (int k = 0; k < 3; ++k) { L.r1[k] = R.m[3 * k]; L.r2[k] = R.m[3 * k + 1]; L.r3[k] = R.m[3 * k + 2]; L.t[k] = t.v[k]; }
    for (int i = 0; i < NPT; ++i)
        for (int k = 0; k < 3; ++k) L.Q[i][k] = Q[i].v[k];
    return err2;
}

// One AbsKernel step (RPP.cpp:229-332): consumes L.Q, produces R, t, the new L.Q and the error.
__device__ __forceinline__ double abs_kernel(const double* ctx, int stride, Lane& L)
{
    double a[3] = {0, 0, 0}, b[3] = {0, 0, 0};
    for (int i = 0; i < NPT; ++i) {
        double fq[3];
        apply_F(ctx, stride, i, L.Q[i], fq);
        const double px = RPPF_C(C_PX + i), py = RPPF_C(C_PY + i);
        for (int k = 0; k < 3; ++k) { a[k] += px * fq[k]; b[k] += py * fq[k]; }
    }
    const double S11 = a[0] * a[0] + a[1] * a[1] + a[2] * a[2], S22 = b[0] * b[0] + b[1] * b[1] + b[2] * b[2],
                 S12 = a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
    const double n0 = a[1] * b[2] - a[2] * b[1], n1 = a[2] * b[0] - a[0] * b[2], n2 = a[0] * b[1] - a[1] * b[0];
    const double s = sqrt(n0 * n0 + n1 * n1 + n2 * n2), tr = S11 + S22;
    if (!(s > 1e-10 * tr)) return abs_kernel_general(ctx, stride, L);
    const double inv = 1.0 / (sqrt(tr + 2.0 * s) * s);
    const double c11 = (S22 + s) * inv, c12 = -S12 * inv, c22 = (S11 + s) * inv;
    double r1[3], r2[3];
    for (int k = 0; k < 3; ++k) { r1[k] = a[k] * c11 + b[k] * c12; r2[k] = a[k] * c12 + b[k] * c22; }
    // t = tF * sum F_i (R P_i)
    double sum[3] = {0, 0, 0}, t[3];
    for (int i = 0; i < NPT; ++i) {
        const double px = RPPF_C(C_PX + i), py = RPPF_C(C_PY + i);
        double rp[3], f[3];
        for (int k = 0; k < 3; ++k) rp[k] = r1[k] * px + r2[k] * py;
        apply_F(ctx, stride, i, rp, f);
        for (int k = 0; k < 3; ++k) sum[k] += f[k];
    }
    mul_tF(ctx, stride, sum, t);
    if (t[2] < 0) return abs_kernel_general(ctx, stride, L);
    for (int k = 0; k < 3; ++k) { L.r1[k] = r1[k]; L.r2[k] = r2[k]; L.t[k] = t[k]; }
    L.r3[0] = r1[1] * r2[2] - r1[2] * r2[1];
    L.r3[1] = r1[2] * r2[0] - r1[0] * r2[2];
    L.r3[2] = r1[0] * r2[1] - r1[1] * r2[0];
    return reproject(ctx, stride, L);
}

// ObjPose with an initial rotation (RPP.cpp:98-128): t from R, Q_i, and the starting error.
__device__ inline double init_from_R(const double* ctx, int stride, const double* R9, Lane& L)
{
    for (int k = 0; k < 3; ++k) { L.r1[k] = R9[3 * k]; L.r2[k] = R9[3 * k + 1]; L.r3[k] = R9[3 * k + 2]; }
    double sum[3] = {0, 0, 0};
    for (int i = 0; i < NPT; ++i) {
        const double px = RPPF_C(C_PX + i), py = RPPF_C(C_PY + i);
        double rp[3], f[3];
        for (int k = 0; k < 3; ++k) rp[k] = L.r1[k] * px + L.r2[k] * py;
        apply_F(ctx, stride, i, rp, f);
        for (int k = 0; k < 3; ++k) sum[k] += f[k] - rp[k];
    }
    mul_tF(ctx, stride, sum, L.t);
    return reproject(ctx, stride, L);
}

// End of ObjPose (RPP.cpp:176-207): R, t (un-centred), object- and image-space errors.
__device__ inline void finalize(const double* ctx, int stride, const Lane& L, double new_err, double* R9, double* t3,
                                double& obj_err, double& img_err)
{
    for (int k = 0; k < 3; ++k) { R9[3 * k] = L.r1[k]; R9[3 * k + 1] = L.r2[k]; R9[3 * k + 2] = L.r3[k]; }
    obj_err = sqrt(new_err / NPT);
    double ie = 0;
    const double u0 = RPPF_C(C_VX + 0), v0 = RPPF_C(C_VY + 0);
    for (int i = 0; i < NPT; ++i) {
        const double px = RPPF_C(C_PX + i), py = RPPF_C(C_PY + i);
        double q[3];
        for (int k = 0; k < 3; ++k) q[k] = (L.r1[k] * px + L.r2[k] * py) + L.t[k];
        // reference quirk (RPP.cpp:197-198): every point is compared with image point 0
        const double xx = q[0] / q[2] - u0, yy = q[1] / q[2] - v0;
        ie += xx * xx + yy * yy;
    }
    img_err = sqrt(ie / NPT);
    const double pbx = RPPF_C(C_PBX), pby = RPPF_C(C_PBY);
    for (int k = 0; k < 3; ++k) t3[k] = L.t[k] - (L.r1[k] * pbx + L.r2[k] * pby);
}

#undef RPPF_C
}  // namespace rppf
