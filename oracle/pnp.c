/*
 * pnp.c -- oracle (TEST INFRASTRUCTURE ONLY, see mantis_oracle.h) for the north star's "Gauss-Newton /
 * PnP pose refinement" in its literal, mantis2 meaning:
 *   computeHypothesis -> cv::solvePnP(obj, img, K, noArray, rvec, tvec, useExtrinsicGuess, SOLVEPNP_ITERATIVE)
 *                                        include/legacy/mantis2/PoseEstimator.h:91-118
 *   computeHypothesisQuadReprojectionError = sqrt(sum |reproj - meas|^2)             :16-25
 *
 * cv::solvePnP(ITERATIVE) is third-party code (OpenCV calib3d, cvFindExtrinsicCameraParams2): for a
 * planar object without a guess it estimates a homography, reads R, t off its columns and then runs
 * Levenberg-Marquardt on the pixel reprojection error; with a guess it runs the same LM from the
 * guess.  This file restates that algorithm (homography by normal equations with h33 = 1 after
 * object-point normalisation, then LM on SE(3) with a left-multiplicative rotation update) -- it does
 * not reproduce OpenCV's iteration path, only its fixed point: the minimiser of the reprojection
 * error in the basin of the homography pose.  Pinned against cv2.solvePnP 4.13.0 on the reference's
 * own vectors (test/legacy/unit_test.cpp:49-72) and on seeded 200-correspondence cases
 * (tests/golden/pnp_golden.json) to 1e-6 rad / 1e-6 m.
 */
#include <math.h>
#include <string.h>

#include "mantis_oracle.h"

static void mat3_mul(const double* a, const double* b, double* r)
{
    double o[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) o[i * 3 + j] = a[i * 3] * b[j] + a[i * 3 + 1] * b[3 + j] + a[i * 3 + 2] * b[6 + j];
    memcpy(r, o, sizeof(o));
}

/* Rodrigues: rotation vector -> matrix */
void mo_rodrigues_exp(const double* w, double* R)
{
    double th = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    double A, B; /* sin(th)/th, (1-cos(th))/th^2 */
    if (th < 1e-8) { A = 1.0 - th * th / 6.0; B = 0.5 - th * th / 24.0; }
    else { A = sin(th) / th; B = (1.0 - cos(th)) / (th * th); }
    double x = w[0], y = w[1], z = w[2];
    R[0] = 1.0 - B * (y * y + z * z); R[1] = B * x * y - A * z;        R[2] = B * x * z + A * y;
    R[3] = B * x * y + A * z;         R[4] = 1.0 - B * (x * x + z * z); R[5] = B * y * z - A * x;
    R[6] = B * x * z - A * y;         R[7] = B * y * z + A * x;         R[8] = 1.0 - B * (x * x + y * y);
}

/* matrix -> rotation vector, the branch structure of cv::Rodrigues (theta in [0, pi]) */
void mo_rodrigues_log(const double* R, double* w)
{
    double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
    double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
    double c = (R[0] + R[4] + R[8] - 1.0) * 0.5;
    c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
    double theta = acos(c);
    if (s < 1e-5) {
        if (c > 0) { w[0] = w[1] = w[2] = 0; return; }
        double t;
        t = (R[0] + 1) * 0.5; rx = sqrt(t > 0 ? t : 0);
        t = (R[4] + 1) * 0.5; ry = sqrt(t > 0 ? t : 0) * (R[1] < 0 ? -1.0 : 1.0);
        t = (R[8] + 1) * 0.5; rz = sqrt(t > 0 ? t : 0) * (R[2] < 0 ? -1.0 : 1.0);
        if (fabs(rx) < fabs(ry) && fabs(rx) < fabs(rz) && (R[5] > 0) != (ry * rz > 0)) rz = -rz;
        double n = sqrt(rx * rx + ry * ry + rz * rz);
        theta /= n;
        w[0] = rx * theta; w[1] = ry * theta; w[2] = rz * theta;
    } else {
        double vth = theta / (2 * s);
        w[0] = rx * vth; w[1] = ry * vth; w[2] = rz * vth;
    }
}

/* in-place Cholesky solve of the n x n SPD system A x = b (A row-major, upper part used); returns 0 if
 * a pivot is not positive */
static int chol_solve(double* A, double* b, int n)
{
    for (int j = 0; j < n; j++) {
        double d = A[j * n + j];
        for (int k = 0; k < j; k++) d -= A[k * n + j] * A[k * n + j];
        if (!(d > 0)) return 0;
        d = sqrt(d);
        A[j * n + j] = d;
        for (int i = j + 1; i < n; i++) {
            double v = A[j * n + i];
            for (int k = 0; k < j; k++) v -= A[k * n + j] * A[k * n + i];
            A[j * n + i] = v / d;
        }
    }
    for (int i = 0; i < n; i++) { /* U^T y = b */
        double v = b[i];
        for (int k = 0; k < i; k++) v -= A[k * n + i] * b[k];
        b[i] = v / A[i * n + i];
    }
    for (int i = n - 1; i >= 0; i--) { /* U x = y */
        double v = b[i];
        for (int k = i + 1; k < n; k++) v -= A[i * n + k] * b[k];
        b[i] = v / A[i * n + i];
    }
    return 1;
}

/* normal equations of the reprojection error at (R, t): A = J^T J (6x6), g = J^T r, cost = |r|^2 */
static double pnp_normal_eq(const float* obj, const float* img, int m, const double* K4, const double* R, const double* t,
                            double* A, double* g)
{
    const double fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
    double cost = 0;
    memset(A, 0, 36 * sizeof(double));
    memset(g, 0, 6 * sizeof(double));
    for (int i = 0; i < m; i++) {
        const double X = obj[3 * i], Y = obj[3 * i + 1], Z = obj[3 * i + 2];
        const double qx = R[0] * X + R[1] * Y + R[2] * Z, qy = R[3] * X + R[4] * Y + R[5] * Z,
                     qz = R[6] * X + R[7] * Y + R[8] * Z; /* R X: the arm of the rotation update */
        const double px = qx + t[0], py = qy + t[1], pz = qz + t[2];
        const double iz = 1.0 / pz;
        const double ru = fx * px * iz + cx - (double)img[2 * i], rv = fy * py * iz + cy - (double)img[2 * i + 1];
        /* d(u,v)/dP */
        const double a0 = fx * iz, a2 = -fx * px * iz * iz, b1 = fy * iz, b2 = -fy * py * iz * iz;
        /* dP = -[q]x dw + dt  ->  J = [d/dP * (-[q]x) | d/dP] */
        double Ju[6], Jv[6];
        /* row u = (a0, 0, a2) * (-[q]x), row v = (0, b1, b2) * (-[q]x);  -[q]x = [[0,qz,-qy],[-qz,0,qx],[qy,-qx,0]] */
        Ju[0] = a2 * qy;            Ju[1] = a0 * qz - a2 * qx;  Ju[2] = -a0 * qy;
        Jv[0] = -b1 * qz + b2 * qy; Jv[1] = -b2 * qx;           Jv[2] = b1 * qx;
        Ju[3] = a0; Ju[4] = 0;  Ju[5] = a2;
        Jv[3] = 0;  Jv[4] = b1; Jv[5] = b2;
        for (int r = 0; r < 6; r++) {
            for (int c = r; c < 6; c++) A[r * 6 + c] += Ju[r] * Ju[c] + Jv[r] * Jv[c];
            g[r] += Ju[r] * ru + Jv[r] * rv;
        }
        cost += ru * ru + rv * rv;
    }
    return cost;
}

/* planar initialisation: homography (object plane z = 0 -> normalised image), then R, t from its columns */
static int pnp_init_planar(const float* obj, const float* img, int m, const double* K4, double* R, double* t)
{
    const double fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
    double sx = 0, sy = 0, sxx = 0, syy = 0;
    for (int i = 0; i < m; i++) {
        sx += obj[3 * i]; sy += obj[3 * i + 1];
        sxx += (double)obj[3 * i] * obj[3 * i]; syy += (double)obj[3 * i + 1] * obj[3 * i + 1];
    }
    const double Xc = sx / m, Yc = sy / m;
    const double var = (sxx / m - Xc * Xc) + (syy / m - Yc * Yc);
    if (!(var > 0)) return 0;
    const double s = 1.0 / sqrt(var);
    double M[64], b[8];
    memset(M, 0, sizeof(M));
    memset(b, 0, sizeof(b));
    for (int i = 0; i < m; i++) {
        const double X = s * (obj[3 * i] - Xc), Y = s * (obj[3 * i + 1] - Yc);
        const double x = ((double)img[2 * i] - cx) / fx, y = ((double)img[2 * i + 1] - cy) / fy;
        const double a1[8] = {X, Y, 1, 0, 0, 0, -x * X, -x * Y}, a2[8] = {0, 0, 0, X, Y, 1, -y * X, -y * Y};
        for (int r = 0; r < 8; r++) {
            for (int c = r; c < 8; c++) M[r * 8 + c] += a1[r] * a1[c] + a2[r] * a2[c];
            b[r] += a1[r] * x + a2[r] * y;
        }
    }
    if (!chol_solve(M, b, 8)) return 0;
    /* H' = [b0 b1 b2; b3 b4 b5; b6 b7 1] on normalised object coordinates; H = H' T */
    double h1[3] = {s * b[0], s * b[3], s * b[6]}, h2[3] = {s * b[1], s * b[4], s * b[7]};
    double h3[3] = {b[2] - Xc * h1[0] - Yc * h2[0], b[5] - Xc * h1[1] - Yc * h2[1], 1.0 - Xc * h1[2] - Yc * h2[2]};
    if (h3[2] < 0) for (int k = 0; k < 3; k++) { h1[k] = -h1[k]; h2[k] = -h2[k]; h3[k] = -h3[k]; }
    const double n1 = sqrt(h1[0] * h1[0] + h1[1] * h1[1] + h1[2] * h1[2]), n2 = sqrt(h2[0] * h2[0] + h2[1] * h2[1] + h2[2] * h2[2]);
    if (!(n1 > 0 && n2 > 0)) return 0;
    const double ts = 2.0 / (n1 + n2);
    double r1[3], r2[3], r3[3];
    for (int k = 0; k < 3; k++) { r1[k] = h1[k] / n1; r2[k] = h2[k] / n2; t[k] = h3[k] * ts; }
    const double d = r1[0] * r2[0] + r1[1] * r2[1] + r1[2] * r2[2];
    for (int k = 0; k < 3; k++) r2[k] -= d * r1[k];
    const double n = sqrt(r2[0] * r2[0] + r2[1] * r2[1] + r2[2] * r2[2]);
    if (!(n > 0)) return 0;
    for (int k = 0; k < 3; k++) r2[k] /= n;
    r3[0] = r1[1] * r2[2] - r1[2] * r2[1]; r3[1] = r1[2] * r2[0] - r1[0] * r2[2]; r3[2] = r1[0] * r2[1] - r1[1] * r2[0];
    for (int k = 0; k < 3; k++) { R[k * 3] = r1[k]; R[k * 3 + 1] = r2[k]; R[k * 3 + 2] = r3[k]; }
    return 1;
}

/* LM from (R, t); returns the number of passes over the correspondences */
static int pnp_refine(const float* obj, const float* img, int m, const double* K4, double* R, double* t, int max_iters,
                      double* cost_out)
{
    double A[36], g[6], lambda = 1e-3;
    double cost = pnp_normal_eq(obj, img, m, K4, R, t, A, g);
    int passes = 1;
    for (int it = 0; it < max_iters; it++) {
        double Ad[36], d[6];
        memcpy(Ad, A, sizeof(A));
        for (int k = 0; k < 6; k++) { Ad[k * 6 + k] += lambda * A[k * 6 + k]; d[k] = -g[k]; }
        if (!chol_solve(Ad, d, 6)) { lambda *= 10; if (lambda > 1e12) break; continue; }
        double dR[9], Rn[9], tn[3], An[36], gn[6];
        mo_rodrigues_exp(d, dR);
        mat3_mul(dR, R, Rn);
        for (int k = 0; k < 3; k++) tn[k] = t[k] + d[3 + k];
        const double cn = pnp_normal_eq(obj, img, m, K4, Rn, tn, An, gn);
        passes++;
        if (cn < cost) {
            double step = 0;
            for (int k = 0; k < 6; k++) step = fmax(step, fabs(d[k]));
            const double gain = cost - cn;
            memcpy(R, Rn, sizeof(Rn)); memcpy(t, tn, sizeof(tn)); memcpy(A, An, sizeof(An)); memcpy(g, gn, sizeof(gn));
            cost = cn;
            lambda = fmax(lambda * 0.1, 1e-15);
            if (step < 1e-9 || gain <= 1e-12 * cost) break;
        } else {
            lambda *= 10;
            if (lambda > 1e12) break;
        }
    }
    *cost_out = cost;
    return passes;
}

/*
 * K: row-major 3x3 float (pinhole, no distortion: the legacy call passes noArray).  guess6: optional
 * rvec (3) + tvec (3) for useExtrinsicGuess = true.  flags bit 0: also refine the second planar
 * candidate (normal mirrored about the line of sight to the object centroid) and keep the lower error.
 * out9: rvec (3), tvec (3), reprojection error sqrt(sum |d|^2), passes, status (1 = ok).
 */
int mo_pnp_iterative(const float* obj, const float* img, int m, const float* K, const double* guess6, int flags,
                     int max_iters, double* out9)
{
    const double K4[4] = {K[0], K[4], K[2], K[5]};
    double R[9], t[3], cost;
    memset(out9, 0, 9 * sizeof(double));
    if (m < 4) return 0;
    if (guess6) {
        mo_rodrigues_exp(guess6, R);
        t[0] = guess6[3]; t[1] = guess6[4]; t[2] = guess6[5];
    } else if (!pnp_init_planar(obj, img, m, K4, R, t)) {
        return 0;
    }
    int passes = pnp_refine(obj, img, m, K4, R, t, max_iters, &cost);
    if (flags & 1) {
        /* second candidate: rotate about the axis perpendicular to the viewing ray c (towards the object
         * centroid) and the plane normal n = R e3 so that n goes to its mirror image about c */
        double cx3[3] = {0, 0, 0};
        for (int i = 0; i < m; i++)
            for (int k = 0; k < 3; k++)
                cx3[k] += (R[k * 3] * obj[3 * i] + R[k * 3 + 1] * obj[3 * i + 1] + R[k * 3 + 2] * obj[3 * i + 2] + t[k]) / m;
        const double cn = sqrt(cx3[0] * cx3[0] + cx3[1] * cx3[1] + cx3[2] * cx3[2]);
        double c[3] = {cx3[0] / cn, cx3[1] / cn, cx3[2] / cn}, n[3] = {R[2], R[5], R[8]};
        double ax[3] = {n[1] * c[2] - n[2] * c[1], n[2] * c[0] - n[0] * c[2], n[0] * c[1] - n[1] * c[0]};
        const double sn = sqrt(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]), cs = n[0] * c[0] + n[1] * c[1] + n[2] * c[2];
        if (sn > 1e-6) {
            const double ang = 2.0 * atan2(sn, cs); /* n -> c is atan2(sn, cs); the mirror image is twice as far */
            double w[3] = {ax[0] / sn * ang, ax[1] / sn * ang, ax[2] / sn * ang}, dR[9], R2[9], t2[3], cost2;
            mo_rodrigues_exp(w, dR);
            mat3_mul(dR, R, R2);
            /* keep the centroid where it is: t2 = centroid - R2 * object centroid */
            double oc[3] = {0, 0, 0};
            for (int i = 0; i < m; i++) for (int k = 0; k < 3; k++) oc[k] += obj[3 * i + k] / (double)m;
            for (int k = 0; k < 3; k++) t2[k] = cx3[k] - (R2[k * 3] * oc[0] + R2[k * 3 + 1] * oc[1] + R2[k * 3 + 2] * oc[2]);
            passes += pnp_refine(obj, img, m, K4, R2, t2, max_iters, &cost2);
            if (cost2 < cost) { memcpy(R, R2, sizeof(R2)); memcpy(t, t2, sizeof(t2)); cost = cost2; }
        }
    }
    mo_rodrigues_log(R, out9);
    out9[3] = t[0]; out9[4] = t[1]; out9[5] = t[2];
    out9[6] = sqrt(cost);
    out9[7] = passes;
    out9[8] = 1;
    return 1;
}
