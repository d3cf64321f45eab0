/*
 * rpp.c -- oracle (test infrastructure) restatement of the Robust Planar Pose estimator the
 * reference runs per quad:
 *   CoPlanarPoseEstimator::estimatePose   include/mantis3/CoPlanarPoseEstimator.cpp:16-58
 *   RPP::Rpp                              include/mantis3/RobustPlanarPose/RPP.cpp:13-64
 *   ObjPose :66-208, AbsKernel :229-332, EstimateT :210-227, Get2ndPose_Exact :693-753,
 *   GetRfor2ndPose_V_Exact :847-945, GetRotationY_wrtT :947-1222, DecomposeR :755-808,
 *   RpyMat :563, RpyAng :592, RpyAng_X :659, GetRotationbyVector :426
 * The quartic's real roots (the reference calls Jenkins-Traub rpoly_ak1, Rpoly.cpp:11) are
 * found here by nested bracketing on the derivative chain -- same roots, different solver.
 *
 * Validation: against the reference's own RPP.cpp/Rpoly.cpp compiled into
 * oracle/_ref/librpp_ref.so (oracle/ref_build/Makefile) and against the known-answer vector
 * of RobustPlanarPose/demo.cpp:17-38, by tests/test_oracle_parts.py.
 */
#include "mantis_oracle.h"
#include <math.h>
#include <string.h>

#define MAXN MO_RPP_MAX_POINTS

/* diagnostics: AbsKernel calls of the successive ObjPose runs of the last mo_rpp() */
int mo_rpp_debug_iters[8];
int mo_rpp_debug_n;

typedef struct { double m[9]; } m3;
typedef struct { double v[3]; } v3;

static m3 m3_mul(m3 a, m3 b)
{
    m3 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            double s = 0;
            for (int k = 0; k < 3; ++k) s += a.m[i * 3 + k] * b.m[k * 3 + j];
            r.m[i * 3 + j] = s;
        }
    return r;
}
static m3 m3_t(m3 a)
{
    m3 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) r.m[i * 3 + j] = a.m[j * 3 + i];
    return r;
}
static v3 m3_v(m3 a, v3 x)
{
    v3 r;
    for (int i = 0; i < 3; ++i) r.v[i] = a.m[i * 3] * x.v[0] + a.m[i * 3 + 1] * x.v[1] + a.m[i * 3 + 2] * x.v[2];
    return r;
}
static m3 m3_eye(void)
{
    m3 r;
    memset(&r, 0, sizeof(r));
    r.m[0] = r.m[4] = r.m[8] = 1;
    return r;
}
static double m3_det(m3 a)
{
    const double* m = a.m;
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]);
}
static m3 m3_inv(m3 a)
{
    const double* m = a.m;
    double d = m3_det(a);
    m3 r;
    r.m[0] = (m[4] * m[8] - m[5] * m[7]) / d;
    r.m[1] = (m[2] * m[7] - m[1] * m[8]) / d;
    r.m[2] = (m[1] * m[5] - m[2] * m[4]) / d;
    r.m[3] = (m[5] * m[6] - m[3] * m[8]) / d;
    r.m[4] = (m[0] * m[8] - m[2] * m[6]) / d;
    r.m[5] = (m[2] * m[3] - m[0] * m[5]) / d;
    r.m[6] = (m[3] * m[7] - m[4] * m[6]) / d;
    r.m[7] = (m[1] * m[6] - m[0] * m[7]) / d;
    r.m[8] = (m[0] * m[4] - m[1] * m[3]) / d;
    return r;
}

/* 3x3 SVD A = U diag(w) V^T, w descending, by one-sided Jacobi on the columns of A.
 * Null directions of U are completed by a cross product (A is rank <= 2 for planar targets). */
static void svd3(m3 A, m3* U, double w[3], m3* V)
{
    double a[9], v[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    memcpy(a, A.m, sizeof(a));
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
        for (int k = 0; k < 3; ++k) V->m[k * 3 + jj] = v[k * 3 + j];
        if (sv[j] > 1e-12 * sv[ord[0]] && sv[j] > 1e-300) {
            for (int k = 0; k < 3; ++k) U->m[k * 3 + jj] = a[k * 3 + j] / sv[j];
        } else {
            w[jj] = 0;
            null_col[jj] = 1;
            for (int k = 0; k < 3; ++k) U->m[k * 3 + jj] = 0;
        }
    }
    for (int jj = 0; jj < 3; ++jj)
        if (null_col[jj]) {
            int a1 = (jj + 1) % 3, a2 = (jj + 2) % 3;
            double x1[3], x2[3];
            for (int k = 0; k < 3; ++k) { x1[k] = U->m[k * 3 + a1]; x2[k] = U->m[k * 3 + a2]; }
            U->m[0 + jj] = x1[1] * x2[2] - x1[2] * x2[1];
            U->m[3 + jj] = x1[2] * x2[0] - x1[0] * x2[2];
            U->m[6 + jj] = x1[0] * x2[1] - x1[1] * x2[0];
        }
}

static double m3_norm2(m3 A) /* spectral norm, RPP::Norm */
{
    m3 U, V;
    double w[3];
    svd3(A, &U, w, &V);
    return w[0];
}

/* RPP::RpyMat(angs): R = Rz(angs[2]) Ry(angs[1]) Rx(angs[0]) */
static m3 rpy_mat(double a0, double a1, double a2)
{
    double cA = cos(a2), sA = sin(a2), cB = cos(a1), sB = sin(a1), cC = cos(a0), sC = sin(a0);
    double cAsB = cA * sB, sAsB = sA * sB;
    m3 R;
    R.m[0] = cA * cB; R.m[1] = cAsB * sC - sA * cC; R.m[2] = cAsB * cC + sA * sC;
    R.m[3] = sA * cB; R.m[4] = sAsB * sC + cA * cC; R.m[5] = sAsB * cC - cA * sC;
    R.m[6] = -sB;     R.m[7] = cB * sC;             R.m[8] = cB * cC;
    return R;
}

static int rpy_ang(m3 R, double ang[3])
{
    const double* r = R.m;
    double sinB = -r[6], cosB = sqrt(r[0] * r[0] + r[3] * r[3]);
    if (fabs(cosB) > 1e-15) {
        ang[0] = atan2(r[7] / cosB, r[8] / cosB);
        ang[1] = atan2(sinB, cosB);
        ang[2] = atan2(r[3] / cosB, r[0] / cosB);
    } else {
        ang[0] = atan2((r[1] - r[5]) / 2, (r[4] + r[2]) / 2);
        ang[1] = M_PI_2;
        ang[2] = 0;
        if (sinB < 0) { ang[0] = -ang[0]; ang[1] = -ang[1]; ang[2] = -ang[2]; }
    }
    m3 d = rpy_mat(ang[0], ang[1], ang[2]);
    for (int i = 0; i < 9; ++i) d.m[i] = R.m[i] - d.m[i];
    return !(m3_norm2(d) > 1e-6);
}

static int rpy_ang_x(m3 R, double ang[3])
{
    if (!rpy_ang(R, ang)) return 0;
    while (fabs(ang[0]) > M_PI_2) {
        if (ang[0] > 0) {
            ang[0] = ang[0] + M_PI; ang[1] = 3 * M_PI - ang[1]; ang[2] = ang[2] + M_PI;
            ang[0] -= 2 * M_PI; ang[1] -= 2 * M_PI; ang[2] -= 2 * M_PI;
        } else {
            ang[0] = ang[0] + M_PI; ang[1] = 3 * M_PI - ang[1]; ang[2] = ang[2] + M_PI;
        }
    }
    return 1;
}

typedef struct {
    int n;
    v3 P[MAXN];   /* centred model points */
    v3 Qp[MAXN];  /* image rays (x, y, 1) */
    m3 F[MAXN];   /* line-of-sight projectors v v^T / v^T v */
    m3 tF;        /* (I - sum F / n)^-1 / n */
    v3 pbar;
} objpose_ctx;

static v3 estimate_t(const objpose_ctx* c, m3 R)
{
    v3 sum = {{0, 0, 0}};
    for (int i = 0; i < c->n; ++i) {
        v3 x = m3_v(c->F[i], m3_v(R, c->P[i]));
        for (int k = 0; k < 3; ++k) sum.v[k] += x.v[k];
    }
    return m3_v(c->tF, sum);
}

/* RPP::AbsKernel.  Q holds the current camera-frame points and is overwritten with the
 * transformed model points (the reference's Qout aliases Q). */
static void abs_kernel(const objpose_ctx* c, v3* Q, m3* R, v3* t, double* err2)
{
    int n = c->n;
    v3 qbar = {{0, 0, 0}};
    for (int i = 0; i < n; ++i) {
        Q[i] = m3_v(c->F[i], Q[i]);
        for (int k = 0; k < 3; ++k) qbar.v[k] += Q[i].v[k];
    }
    for (int k = 0; k < 3; ++k) qbar.v[k] /= n;
    (void)qbar; /* the reference computes qbar but does not subtract it (RPP.cpp:246-254) */
    m3 M;
    memset(&M, 0, sizeof(M));
    for (int i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) M.m[a * 3 + b] += c->P[i].v[a] * Q[i].v[b];
    m3 U, V;
    double w[3];
    svd3(M, &U, w, &V);
    m3 Ut = m3_t(U);
    *R = m3_mul(V, Ut);
    double d = m3_det(*R);
    if (d > 0) {
        *t = estimate_t(c, *R);
        if (t->v[2] < 0) {
            for (int k = 0; k < 3; ++k) V.m[k * 3 + 2] = -V.m[k * 3 + 2];
            *R = m3_mul(V, Ut);
            for (int k = 0; k < 9; ++k) R->m[k] = -R->m[k];
            *t = estimate_t(c, *R);
        }
    } else {
        for (int k = 0; k < 3; ++k) V.m[k * 3 + 2] = -V.m[k * 3 + 2];
        *R = m3_mul(V, Ut);
        *t = estimate_t(c, *R);
        if (t->v[2] < 0) {
            for (int k = 0; k < 9; ++k) R->m[k] = -R->m[k];
            *t = estimate_t(c, *R);
        }
    }
    *err2 = 0;
    for (int i = 0; i < n; ++i) {
        v3 q = m3_v(*R, c->P[i]);
        for (int k = 0; k < 3; ++k) q.v[k] += t->v[k];
        Q[i] = q;
        v3 fq = m3_v(c->F[i], q);
        for (int k = 0; k < 3; ++k) { double e = q.v[k] - fq.v[k]; *err2 += e * e; }
    }
}

static void objpose_setup(objpose_ctx* c, const double* model3xn, const double* iprts3xn, int n)
{
    c->n = n;
    for (int k = 0; k < 3; ++k) {
        double s = 0;
        for (int i = 0; i < n; ++i) s += model3xn[k * n + i];
        c->pbar.v[k] = s / n;
    }
    m3 sumF;
    memset(&sumF, 0, sizeof(sumF));
    for (int i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) {
            c->P[i].v[k] = model3xn[k * n + i] - c->pbar.v[k];
            c->Qp[i].v[k] = iprts3xn[k * n + i];
        }
        const double* v = c->Qp[i].v;
        double vv = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) {
                c->F[i].m[a * 3 + b] = (v[a] * v[b]) / vv;
                sumF.m[a * 3 + b] += c->F[i].m[a * 3 + b];
            }
    }
    m3 A = m3_eye();
    for (int k = 0; k < 9; ++k) A.m[k] -= sumF.m[k] / n;
    c->tF = m3_inv(A);
    for (int k = 0; k < 9; ++k) c->tF.m[k] /= n;
}

/* RPP::ObjPose.  has_init: start from initR, else weak-perspective start. */
static void obj_pose(const objpose_ctx* c, int has_init, m3 initR, m3* Rout, v3* tout, int* it, double* obj_err,
                     double* img_err)
{
    const double TOL = 1e-5, EPSILON = 1e-8;
    int n = c->n;
    v3 Q[MAXN];
    m3 Ri;
    v3 ti;
    double old_err, new_err;
    *it = 0;
    if (has_init) {
        Ri = initR;
        v3 sum = {{0, 0, 0}};
        for (int i = 0; i < n; ++i) {
            v3 rp = m3_v(Ri, c->P[i]);
            v3 frp = m3_v(c->F[i], rp);
            for (int k = 0; k < 3; ++k) sum.v[k] += frp.v[k] - rp.v[k];
        }
        ti = m3_v(c->tF, sum);
        old_err = 0;
        for (int i = 0; i < n; ++i) {
            v3 q = m3_v(Ri, c->P[i]);
            for (int k = 0; k < 3; ++k) q.v[k] += ti.v[k];
            Q[i] = q;
            v3 fq = m3_v(c->F[i], q);
            for (int k = 0; k < 3; ++k) { double e = q.v[k] - fq.v[k]; old_err += e * e; }
        }
    } else {
        for (int i = 0; i < n; ++i) Q[i] = c->Qp[i];
        abs_kernel(c, Q, &Ri, &ti, &old_err);
        *it = 1;
    }
    abs_kernel(c, Q, &Ri, &ti, &new_err);
    *it += 1;
    while (fabs((old_err - new_err) / old_err) > TOL && new_err > EPSILON) {
        old_err = new_err;
        abs_kernel(c, Q, &Ri, &ti, &new_err);
        *it += 1;
        if (*it > 100000) break; /* the reference has no cap; guard the oracle against NaN loops */
    }
    if (mo_rpp_debug_n < 8) mo_rpp_debug_iters[mo_rpp_debug_n++] = *it;
    *Rout = Ri;
    *obj_err = sqrt(new_err / n);
    double ie = 0;
    for (int i = 0; i < n; ++i) {
        v3 q = m3_v(Ri, c->P[i]);
        for (int k = 0; k < 3; ++k) q.v[k] += ti.v[k];
        /* reference quirk: every point is compared with image point 0 (RPP.cpp:197-198) */
        double xx = q.v[0] / q.v[2] - c->Qp[0].v[0];
        double yy = q.v[1] / q.v[2] - c->Qp[0].v[1];
        ie += xx * xx + yy * yy;
    }
    *img_err = sqrt(ie / n);
    v3 rp = m3_v(Ri, c->pbar);
    for (int k = 0; k < 3; ++k) tout->v[k] = ti.v[k] - rp.v[k];
}

/* ---- real roots of a polynomial of degree <= 4 (coefficients descending) ---- */
static double poly_eval(const double* c, int d, double x)
{
    double y = c[0];
    for (int i = 1; i <= d; ++i) y = y * x + c[i];
    return y;
}

static int poly_real_roots(const double* c_in, int d, double* roots)
{
    const double* c = c_in;
    while (d > 0 && c[0] == 0.0) { ++c; --d; }
    if (d <= 0) return 0;
    if (d == 1) { roots[0] = -c[1] / c[0]; return 1; }
    if (d == 2) {
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
    double dc[4], crit[4];
    for (int i = 0; i < d; ++i) dc[i] = c[i] * (d - i);
    int nc = poly_real_roots(dc, d - 1, crit);
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

/* RPP::GetRotationY_wrtT: stationary points of the object-space error as a function of the
 * rotation beta about y; returns the minima (angles in radians) and the optimal t for each. */
static int rotation_y_wrt_t(const v3* v, const v3* p, int n, m3 Rz, double* bl, v3* tnew)
{
    m3 V[MAXN], G, I = m3_eye();
    memset(&G, 0, sizeof(G));
    for (int i = 0; i < n; ++i) {
        double vv = v[i].v[0] * v[i].v[0] + v[i].v[1] * v[i].v[1] + v[i].v[2] * v[i].v[2];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) {
                V[i].m[a * 3 + b] = (v[i].v[a] * v[i].v[b]) / vv;
                G.m[a * 3 + b] += V[i].m[a * 3 + b];
            }
    }
    for (int k = 0; k < 9; ++k) G.m[k] = I.m[k] - G.m[k] / n;
    G = m3_inv(G);
    for (int k = 0; k < 9; ++k) G.m[k] /= n;

    m3 opt_t;
    memset(&opt_t, 0, sizeof(opt_t));
    m3 Rpi[MAXN];
    for (int i = 0; i < n; ++i) {
        double px = p[i].v[0], py = p[i].v[1], pz = p[i].v[2];
        double r[9] = {-px, 2 * pz, px, py, 0, py, -pz, -2 * px, pz};
        memcpy(Rpi[i].m, r, sizeof(r));
        m3 A;
        for (int k = 0; k < 9; ++k) A.m[k] = V[i].m[k] - I.m[k];
        m3 B = m3_mul(m3_mul(A, Rz), Rpi[i]);
        for (int k = 0; k < 9; ++k) opt_t.m[k] += B.m[k];
    }
    opt_t = m3_mul(G, opt_t);

    double E2[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < n; ++i) {
        m3 A, S = m3_mul(Rz, Rpi[i]);
        for (int k = 0; k < 9; ++k) { A.m[k] = I.m[k] - V[i].m[k]; S.m[k] += opt_t.m[k]; }
        m3 E = m3_mul(A, S);
        for (int k = 0; k < 3; ++k) {
            double e2 = E.m[k * 3 + 0], e1 = E.m[k * 3 + 1], e0 = E.m[k * 3 + 2];
            E2[0] += e2 * e2;
            E2[1] += 2 * e1 * e2;
            E2[2] += 2 * e0 * e2 + e1 * e1;
            E2[3] += 2 * e0 * e1;
            E2[4] += e0 * e0;
        }
    }
    double e4 = E2[0], e3 = E2[1], e2 = E2[2], e1 = E2[3], e0 = E2[4];
    double a4 = -e3, a3 = 4 * e4 - 2 * e2, a2 = -3 * e1 + 3 * e3, a1 = -4 * e0 + 2 * e2, a0 = e1;
    double coeffs[5] = {a4, a3, a2, a1, a0}, roots[4];
    int nr = poly_real_roots(coeffs, 4, roots);
    int nsol = 0;
    for (int i = 0; i < nr; ++i) {
        double at = roots[i];
        if (!(fabs(pow(1.0 + at * at, 3.0)) > 0.1)) continue;
        double sa = (2.0 * at) / (1.0 + at * at), ca = (1.0 - at * at) / (1.0 + at * at);
        double al_deg = atan2(sa, ca) * 180 / M_PI;
        double tmm = 4 * a4 * at * at * at + 3 * a3 * at * at + 2 * a2 * at + a1;
        if (!(tmm > 0)) continue;
        double beta = al_deg * M_PI / 180; /* the reference round-trips through degrees */
        m3 R = m3_mul(Rz, rpy_mat(0, beta, 0));
        v3 topt = {{0, 0, 0}};
        for (int k = 0; k < n; ++k) {
            v3 rp = m3_v(R, p[k]);
            v3 vrp = m3_v(V[k], rp);
            for (int q = 0; q < 3; ++q) topt.v[q] += vrp.v[q] - rp.v[q];
        }
        bl[nsol] = (al_deg / 180) * M_PI;
        tnew[nsol] = m3_v(G, topt);
        nsol++;
    }
    return nsol;
}

/* RPP::Get2ndPose_Exact + GetRfor2ndPose_V_Exact + DecomposeR: candidate second poses */
static int second_poses(const double* model3xn, const double* iprts3xn, int n, m3 R, v3 t, m3* solR, v3* solT)
{
    /* cent = normalised sum of the unit image rays */
    v3 cent = {{0, 0, 0}};
    for (int i = 0; i < n; ++i) {
        double x = iprts3xn[i], y = iprts3xn[n + i], z = iprts3xn[2 * n + i];
        double m = 1.0 / sqrt(x * x + y * y + z * z);
        cent.v[0] += x * m; cent.v[1] += y * m; cent.v[2] += z * m;
    }
    for (int k = 0; k < 3; ++k) cent.v[k] /= 3; /* RPP::Mean divides by cols (=3); harmless before NormRv */
    {
        double m = 1.0 / sqrt(cent.v[0] * cent.v[0] + cent.v[1] * cent.v[1] + cent.v[2] * cent.v[2]);
        for (int k = 0; k < 3; ++k) cent.v[k] *= m;
    }
    /* Rim = GetRotationbyVector((0,0,1), cent): rotation by acos(cent.z) about cent x z */
    double winkel = acos(cent.v[2]);
    double ax = cent.v[1] * 1.0 - cent.v[2] * 0.0, ay = cent.v[2] * 0.0 - cent.v[0] * 1.0, az = 0.0;
    double an = sqrt(ax * ax + ay * ay + az * az);
    double sh = sin(winkel * 0.5), qa = cos(winkel * 0.5);
    double qb = ax / an * sh, qc = ay / an * sh, qd = az / an * sh;
    double qn = 1.0 / sqrt(qa * qa + qb * qb + qc * qc + qd * qd);
    qa *= qn; qb *= qn; qc *= qn; qd *= qn;
    m3 Rim;
    Rim.m[0] = qa * qa + qb * qb - qc * qc - qd * qd; Rim.m[1] = 2 * (qb * qc - qa * qd); Rim.m[2] = 2 * (qb * qd + qa * qc);
    Rim.m[3] = 2 * (qb * qc + qa * qd); Rim.m[4] = qa * qa + qc * qc - qb * qb - qd * qd; Rim.m[5] = 2 * (qc * qd - qa * qb);
    Rim.m[6] = 2 * (qb * qd - qa * qc); Rim.m[7] = 2 * (qc * qd + qa * qb); Rim.m[8] = qa * qa + qd * qd - qb * qb - qc * qc;

    v3 v_[MAXN], P[MAXN];
    for (int i = 0; i < n; ++i) {
        v3 vi = {{iprts3xn[i], iprts3xn[n + i], iprts3xn[2 * n + i]}};
        v_[i] = m3_v(Rim, vi);
        P[i].v[0] = model3xn[i]; P[i].v[1] = model3xn[n + i]; P[i].v[2] = model3xn[2 * n + i];
    }
    m3 R_ = m3_mul(Rim, R);
    v3 t_ = m3_v(Rim, t);
    (void)t_;

    /* DecomposeR */
    double cl = atan2(R_.m[7], R_.m[6]);
    m3 RzN = rpy_mat(0, 0, cl);
    m3 Rt = m3_mul(R_, RzN);
    double ang[3];
    if (Rt.m[7] > 1e-3) return -1;
    if (!rpy_ang_x(Rt, ang)) return -1;
    if (fabs(ang[0]) > 1e-3) return -1;
    m3 Rz2 = m3_mul(RzN, rpy_mat(0, 0, M_PI));
    m3 Rt2 = m3_mul(R_, Rz2);
    if (Rt2.m[7] > 1e-3) return -1;
    if (!rpy_ang_x(Rt2, ang)) return -1;

    m3 RzNt = m3_t(RzN);
    v3 P_[MAXN];
    for (int i = 0; i < n; ++i) P_[i] = m3_v(RzNt, P[i]);
    if (!rpy_ang_x(Rt, ang)) return -1;
    m3 Rz = rpy_mat(0, 0, ang[2]);

    double bl[4];
    v3 tnew[4];
    int ns = rotation_y_wrt_t(v_, P_, n, Rz, bl, tnew);
    if (ns == 0) return -1;
    m3 Rimt = m3_t(Rim);
    for (int j = 0; j < ns; ++j) {
        m3 Rs = m3_mul(m3_mul(Rz, rpy_mat(0, bl[j], 0)), RzNt);
        solR[j] = m3_mul(Rimt, Rs);
        solT[j] = m3_v(Rimt, tnew[j]);
    }
    return ns;
}

/* RPP::Rpp.  model / iprts are 3 x n row-major (as the reference's cv::Mat).  Returns 1 on
 * success (0 where the reference returns false). */
int mo_rpp(const double* model3xn, const double* iprts3xn, int n, double* R9, double* t3, int* iters,
           double* obj_err, double* img_err)
{
    if (n < 3 || n > MAXN) return 0;
    mo_rpp_debug_n = 0;
    objpose_ctx c;
    objpose_setup(&c, model3xn, iprts3xn, n);
    m3 Rlu;
    v3 tlu;
    int it;
    double oe, ie;
    obj_pose(&c, 0, m3_eye(), &Rlu, &tlu, &it, &oe, &ie);
    m3 solR[4];
    v3 solT[4];
    int ns = second_poses(model3xn, iprts3xn, n, Rlu, tlu, solR, solT);
    if (ns <= 0) return 0;
    int best = -1;
    double lowest = 1e6, boe = 0, bie = 0;
    m3 bR = Rlu;
    v3 bt = tlu;
    for (int i = 0; i < ns; ++i) {
        m3 Ri;
        v3 ti;
        obj_pose(&c, 1, solR[i], &Ri, &ti, &it, &oe, &ie);
        if (oe < lowest) { lowest = oe; best = i; bR = Ri; bt = ti; boe = oe; bie = ie; }
    }
    if (best < 0) return 0; /* the reference indexes sol[-1] here (undefined behaviour) */
    memcpy(R9, bR.m, sizeof(double) * 9);
    memcpy(t3, bt.v, sizeof(double) * 3);
    *iters = it;
    *obj_err = boe;
    *img_err = bie;
    return 1;
}
