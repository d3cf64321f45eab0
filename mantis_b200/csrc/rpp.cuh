// rpp.cuh -- device implementation of the Robust Planar Pose estimator the reference runs per quad
//   CoPlanarPoseEstimator::estimatePose   include/mantis3/CoPlanarPoseEstimator.cpp:16-58
//   RPP::Rpp  RPP.cpp:13-64;  ObjPose :66-208;  AbsKernel :229-332;  EstimateT :210-227;
//   Get2ndPose_Exact :693-753;  GetRfor2ndPose_V_Exact :847-945;  GetRotationY_wrtT :947-1222;
//   DecomposeR :755-808;  RpyMat :563;  RpyAng :592;  RpyAng_X :659
// One thread per quad, fp64, everything in registers / local memory: orthogonal iteration with a
// 3x3 one-sided Jacobi SVD per step, then the second-pose search (quartic in tan(beta/2), real
// roots by nested bracketing instead of the reference's Jenkins-Traub) and a refinement from
// every candidate; the lowest object-space error wins.
#pragma once
#include "posemath.cuh"

namespace rpp {
using pm::M3;
using pm::V3;

// A = U diag(w) V^T, w descending; null directions of U completed by a cross product.
__device__ inline void svd3(const M3& A, M3& U, double w[3], M3& V)
{
    double a[9], v[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    for (int i = 0; i < 9; ++i) a[i] = A.m[i];
    for (int sweep = 0; sweep < 60; ++sweep) {
        int changed = 0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double al = 0, be = 0, ga = 0;
                for (int k = 0; k < 3; ++k) {
                    al += a[k * 3 + p] * a[k * 3 + p];
                    be += a[k * 3 + q] * a[k * 3 + q];
                    ga += a[k * 3 + p] * a[k * 3 + q];
                }
                if (fabs(ga) <= 1e-300 || fabs(ga) <= 2.220446049250313e-16 * sqrt(al * be)) continue;
                if (al <= 1e-28 * be || be <= 1e-28 * al) continue; /* a numerically null column is already orthogonal */
                changed = 1;
                double zeta = (be - al) / (2.0 * ga);
                double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int k = 0; k < 3; ++k) {
                    double ap = a[k * 3 + p], aq = a[k * 3 + q];
                    a[k * 3 + p] = c * ap - s * aq;
                    a[k * 3 + q] = s * ap + c * aq;
                    double vp = v[k * 3 + p], vq = v[k * 3 + q];
                    v[k * 3 + p] = c * vp - s * vq;
                    v[k * 3 + q] = s * vp + c * vq;
                }
            }
        if (!changed) break;
    }
    double sv[3];
    int ord[3] = {0, 1, 2};
    for (int j = 0; j < 3; ++j) sv[j] = sqrt(a[j] * a[j] + a[3 + j] * a[3 + j] + a[6 + j] * a[6 + j]);
    for (int i = 0; i < 3; ++i)
        for (int j = i + 1; j < 3; ++j)
            if (sv[ord[j]] > sv[ord[i]]) { int t = ord[i]; ord[i] = ord[j]; ord[j] = t; }
    int null_col[3] = {0, 0, 0};
    for (int jj = 0; jj < 3; ++jj) {
        int j = ord[jj];
        w[jj] = sv[j];
        for (int k = 0; k < 3; ++k) V.m[k * 3 + jj] = v[k * 3 + j];
        if (sv[j] > 1e-12 * sv[ord[0]] && sv[j] > 1e-300) {
            for (int k = 0; k < 3; ++k) U.m[k * 3 + jj] = a[k * 3 + j] / sv[j];
        } else {
            w[jj] = 0;
            null_col[jj] = 1;
            for (int k = 0; k < 3; ++k) U.m[k * 3 + jj] = 0;
        }
    }
    for (int jj = 0; jj < 3; ++jj)
        if (null_col[jj]) {
            int a1 = (jj + 1) % 3, a2 = (jj + 2) % 3;
            double x1[3], x2[3];
            for (int k = 0; k < 3; ++k) { x1[k] = U.m[k * 3 + a1]; x2[k] = U.m[k * 3 + a2]; }
            U.m[0 + jj] = x1[1] * x2[2] - x1[2] * x2[1];
            U.m[3 + jj] = x1[2] * x2[0] - x1[0] * x2[2];
            U.m[6 + jj] = x1[0] * x2[1] - x1[1] * x2[0];
        }
}

__device__ inline double norm2(const M3& A)
{
    M3 U, V;
    double w[3];
    svd3(A, U, w, V);
    return w[0];
}

__device__ inline M3 rpy_mat(double a0, double a1, double a2)
{
    double cA = cos(a2), sA = sin(a2), cB = cos(a1), sB = sin(a1), cC = cos(a0), sC = sin(a0);
    double cAsB = cA * sB, sAsB = sA * sB;
    M3 R;
    R.m[0] = cA * cB; R.m[1] = cAsB * sC - sA * cC; R.m[2] = cAsB * cC + sA * sC;
    R.m[3] = sA * cB; R.m[4] = sAsB * sC + cA * cC; R.m[5] = sAsB * cC - cA * sC;
    R.m[6] = -sB;     R.m[7] = cB * sC;             R.m[8] = cB * cC;
    return R;
}

__device__ inline bool rpy_ang(const M3& R, double ang[3])
{
    const double PI_2 = 1.57079632679489661923;
    const double* r = R.m;
    double sinB = -r[6], cosB = sqrt(r[0] * r[0] + r[3] * r[3]);
    if (fabs(cosB) > 1e-15) {
        ang[0] = atan2(r[7] / cosB, r[8] / cosB);
        ang[1] = atan2(sinB, cosB);
        ang[2] = atan2(r[3] / cosB, r[0] / cosB);
    } else {
        ang[0] = atan2((r[1] - r[5]) / 2, (r[4] + r[2]) / 2);
        ang[1] = PI_2;
        ang[2] = 0;
        if (sinB < 0) { ang[0] = -ang[0]; ang[1] = -ang[1]; ang[2] = -ang[2]; }
    }
    M3 d = rpy_mat(ang[0], ang[1], ang[2]);
    for (int i = 0; i < 9; ++i) d.m[i] = R.m[i] - d.m[i];
    return !(norm2(d) > 1e-6);
}

__device__ inline bool rpy_ang_x(const M3& R, double ang[3])
{
    const double PI = 3.14159265358979323846, PI_2 = 1.57079632679489661923;
    if (!rpy_ang(R, ang)) return false;
    int guard = 0;
    while (fabs(ang[0]) > PI_2 && guard++ < 64) {
        if (ang[0] > 0) {
            ang[0] = ang[0] + PI; ang[1] = 3 * PI - ang[1]; ang[2] = ang[2] + PI;
            ang[0] -= 2 * PI; ang[1] -= 2 * PI; ang[2] -= 2 * PI;
        } else {
            ang[0] = ang[0] + PI; ang[1] = 3 * PI - ang[1]; ang[2] = ang[2] + PI;
        }
    }
    return true;
}

template <int N>
struct ObjCtx {
    V3 P[N];   // centred model points
    V3 Qp[N];  // image rays (x, y, 1)
    M3 F[N];   // v v^T / v^T v
    M3 tF;     // (I - sum F / n)^-1 / n
    V3 pbar;
};

template <int N>
__device__ inline V3 estimate_t(const ObjCtx<N>& c, const M3& R, int n = N)
{
    V3 sum = {{0, 0, 0}};
    for (int i = 0; i < n; ++i) {
        V3 x = pm::mulv(c.F[i], pm::mulv(R, c.P[i]));
        for (int k = 0; k < 3; ++k) sum.v[k] += x.v[k];
    }
    return pm::mulv(c.tF, sum);
}

template <int N>
__device__ inline void abs_kernel(const ObjCtx<N>& c, V3* Q, M3& R, V3& t, double& err2, int n = N)
{
    for (int i = 0; i < n; ++i) Q[i] = pm::mulv(c.F[i], Q[i]);
    M3 M;
    for (int k = 0; k < 9; ++k) M.m[k] = 0;
    for (int i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) M.m[a * 3 + b] += c.P[i].v[a] * Q[i].v[b];
    M3 U, V;
    double w[3];
    svd3(M, U, w, V);
    M3 Ut = pm::transpose(U);
    R = pm::mul(V, Ut);
    if (pm::det(R) > 0) {
        t = estimate_t(c, R, n);
        if (t.v[2] < 0) {
            for (int k = 0; k < 3; ++k) V.m[k * 3 + 2] = -V.m[k * 3 + 2];
            R = pm::mul(V, Ut);
            for (int k = 0; k < 9; ++k) R.m[k] = -R.m[k];
            t = estimate_t(c, R, n);
        }
    } else {
        for (int k = 0; k < 3; ++k) V.m[k * 3 + 2] = -V.m[k * 3 + 2];
        R = pm::mul(V, Ut);
        t = estimate_t(c, R, n);
        if (t.v[2] < 0) {
            for (int k = 0; k < 9; ++k) R.m[k] = -R.m[k];
            t = estimate_t(c, R, n);
        }
    }
    err2 = 0;
    for (int i = 0; i < n; ++i) {
        V3 q = pm::mulv(R, c.P[i]);
        for (int k = 0; k < 3; ++k) q.v[k] += t.v[k];
        Q[i] = q;
        V3 fq = pm::mulv(c.F[i], q);
        for (int k = 0; k < 3; ++k) { double e = q.v[k] - fq.v[k]; err2 += e * e; }
    }
}

// model / rays: N points, model[i] = (x, y, z), ray[i] = (x, y, 1)
template <int N>
__device__ inline void setup(ObjCtx<N>& c, const V3* model, const V3* ray, int n = N)
{
    for (int k = 0; k < 3; ++k) {
        double s = 0;
        for (int i = 0; i < n; ++i) s += model[i].v[k];
        c.pbar.v[k] = s / n;
    }
    M3 sumF;
    for (int k = 0; k < 9; ++k) sumF.m[k] = 0;
    for (int i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) {
            c.P[i].v[k] = model[i].v[k] - c.pbar.v[k];
            c.Qp[i].v[k] = ray[i].v[k];
        }
        const double* v = c.Qp[i].v;
        double vv = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) {
                c.F[i].m[a * 3 + b] = (v[a] * v[b]) / vv;
                sumF.m[a * 3 + b] += c.F[i].m[a * 3 + b];
            }
    }
    M3 A = pm::eye();
    for (int k = 0; k < 9; ++k) A.m[k] -= sumF.m[k] / n;
    c.tF = pm::inverse(A);
    for (int k = 0; k < 9; ++k) c.tF.m[k] /= n;
}

template <int N>
__device__ inline void obj_pose(const ObjCtx<N>& c, bool has_init, const M3& initR, M3& Rout, V3& tout, double& obj_err,
                                double& img_err, int* calls = nullptr, int n = N)
{
    const double TOL = 1e-5, EPSILON = 1e-8;
    V3 Q[N];
    M3 Ri;
    V3 ti;
    double old_err, new_err;
    int it = 0;
    if (has_init) {
        Ri = initR;
        V3 sum = {{0, 0, 0}};
        for (int i = 0; i < n; ++i) {
            V3 rp = pm::mulv(Ri, c.P[i]);
            V3 frp = pm::mulv(c.F[i], rp);
            for (int k = 0; k < 3; ++k) sum.v[k] += frp.v[k] - rp.v[k];
        }
        ti = pm::mulv(c.tF, sum);
        old_err = 0;
        for (int i = 0; i < n; ++i) {
            V3 q = pm::mulv(Ri, c.P[i]);
            for (int k = 0; k < 3; ++k) q.v[k] += ti.v[k];
            Q[i] = q;
            V3 fq = pm::mulv(c.F[i], q);
            for (int k = 0; k < 3; ++k) { double e = q.v[k] - fq.v[k]; old_err += e * e; }
        }
    } else {
        for (int i = 0; i < n; ++i) Q[i] = c.Qp[i];
        abs_kernel(c, Q, Ri, ti, old_err, n);
        it = 1;
    }
    abs_kernel(c, Q, Ri, ti, new_err, n);
    it += 1;
    while (fabs((old_err - new_err) / old_err) > TOL && new_err > EPSILON) {
        old_err = new_err;
        abs_kernel(c, Q, Ri, ti, new_err, n);
        if (++it > 100000) break;
    }
    if (calls) *calls = it;
    Rout = Ri;
    obj_err = sqrt(new_err / n);
    double ie = 0;
    for (int i = 0; i < n; ++i) {
        V3 q = pm::mulv(Ri, c.P[i]);
        for (int k = 0; k < 3; ++k) q.v[k] += ti.v[k];
        // reference quirk (RPP.cpp:197-198): every point is compared with image point 0
        double xx = q.v[0] / q.v[2] - c.Qp[0].v[0];
        double yy = q.v[1] / q.v[2] - c.Qp[0].v[1];
        ie += xx * xx + yy * yy;
    }
    img_err = sqrt(ie / n);
    V3 rp = pm::mulv(Ri, c.pbar);
    for (int k = 0; k < 3; ++k) tout.v[k] = ti.v[k] - rp.v[k];
}

// ---- real roots of a polynomial of degree <= 4 (coefficients descending) ----
__device__ inline double poly_eval(const double* c, int d, double x)
{
    double y = c[0];
    for (int i = 1; i <= d; ++i) y = y * x + c[i];
    return y;
}

__device__ inline int quad_roots(const double* c, double* roots)
{
    double disc = c[1] * c[1] - 4 * c[0] * c[2];
    if (disc < 0) return 0;
    double sq = sqrt(disc);
    double q = -0.5 * (c[1] + (c[1] >= 0 ? sq : -sq));
    double r1 = q / c[0], r2 = (q != 0.0) ? c[2] / q : r1;
    if (r1 > r2) { double t = r1; r1 = r2; r2 = t; }
    roots[0] = r1;
    if (disc == 0) return 1;
    roots[1] = r2;
    return 2;
}

// roots of c (degree d >= 3) given the real roots of its derivative
__device__ inline int bracket_roots(const double* c, int d, const double* crit, int nc, double* roots)
{
    double B = 0;
    for (int i = 1; i <= d; ++i) B = fmax(B, fabs(c[i] / c[0]));
    B += 1.0;
    double knots[6];
    int nk = 0;
    knots[nk++] = -B;
    for (int i = 0; i < nc; ++i)
        if (crit[i] > -B && crit[i] < B) knots[nk++] = crit[i];
    knots[nk++] = B;
    int nr = 0;
    for (int i = 0; i + 1 < nk; ++i) {
        double lo = knots[i], hi = knots[i + 1];
        double flo = poly_eval(c, d, lo), fhi = poly_eval(c, d, hi);
        if (flo == 0.0) { if (nr == 0 || roots[nr - 1] != lo) roots[nr++] = lo; continue; }
        if (fhi == 0.0) { roots[nr++] = hi; continue; }
        if ((flo < 0) == (fhi < 0)) continue;
        for (int it = 0; it < 200; ++it) {
            double mid = 0.5 * (lo + hi);
            if (mid <= lo || mid >= hi) break;
            double fm = poly_eval(c, d, mid);
            if (fm == 0.0) { lo = hi = mid; break; }
            if ((fm < 0) == (flo < 0)) { lo = mid; flo = fm; } else { hi = mid; }
        }
        roots[nr++] = 0.5 * (lo + hi);
    }
    return nr;
}

__device__ inline int poly_real_roots(const double* c_in, int d, double* roots)
{
    const double* c = c_in;
    while (d > 0 && c[0] == 0.0) { ++c; --d; }
    if (d <= 0) return 0;
    if (d == 1) { roots[0] = -c[1] / c[0]; return 1; }
    if (d == 2) return quad_roots(c, roots);
    // derivative chain without recursion
    double d1[4], r1[4];
    for (int i = 0; i < d; ++i) d1[i] = c[i] * (d - i);
    int n1;
    if (d == 3) {
        n1 = quad_roots(d1, r1);
    } else {  // d == 4: d1 is a cubic (leading coefficient non-zero because c[0] is)
        double d2[3], r2[2];
        for (int i = 0; i < 3; ++i) d2[i] = d1[i] * (3 - i);
        int n2 = quad_roots(d2, r2);
        n1 = bracket_roots(d1, 3, r2, n2, r1);
    }
    return bracket_roots(c, d, r1, n1, roots);
}

}  // namespace rpp

#include "rpp_second.inl"
