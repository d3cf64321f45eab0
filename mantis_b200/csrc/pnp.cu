// pnp.cu -- K11: batched Levenberg-Marquardt / Gauss-Newton PnP, one warp per frame.
//
// The north star's "Gauss-Newton/PnP pose refinement" in its literal (mantis2) meaning:
//   computeHypothesis -> cv::solvePnP(obj, img, K, noArray, rvec, tvec, guess, SOLVEPNP_ITERATIVE)
//                                          include/legacy/mantis2/PoseEstimator.h:91-118
//   computeHypothesisQuadReprojectionError :16-25  (sqrt of the summed squared pixel residuals)
// Same algorithm as oracle/pnp.c: planar initialisation from a homography (8x8 normal equations,
// h33 = 1, object points normalised), then LM on SE(3) with a left-multiplicative rotation update.
//
// Work split: the lanes of a warp own the correspondences (lane, lane+32, ...); each pass accumulates
// the 21 + 6 + 1 entries of J^T J, J^T r and the cost per lane in registers, a 5-step transposing
// shuffle reduction leaves entry e summed on lane e (31 shuffles of doubles instead of 28 x 5), the
// sums are exchanged through 256 bytes of shared memory, and every lane then factors the damped 6x6
// system with an unrolled Cholesky held entirely in registers -- no lane diverges, no global traffic
// besides re-reading the correspondences (L1-resident: 4 KB per frame).  fp64 throughout: the systems
// are small and the tolerance (1e-4 rad, 1 mm against OpenCV) leaves no room for fp32 normal equations.
#include <math.h>
#include "ctx.h"

namespace {

constexpr int PNT = 128;  // 4 warps = 4 frames per CTA

// Transposing reduction: v[0..31] per lane in, v[0] = sum over lanes of entry `lane` out.
__device__ __forceinline__ void warp_transpose_sum32(double (&v)[32], int lane)
{
#pragma unroll
    for (int c = 32, o = 16; o >= 1; c >>= 1, o >>= 1) {
        const bool hi = lane & o;
#pragma unroll
        for (int k = 0; k < c / 2; ++k) {
            const double send = hi ? v[k] : v[k + c / 2];
            const double keep = hi ? v[k + c / 2] : v[k];
            v[k] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
}

// Cholesky solve of an N x N SPD system held in registers (upper triangle of A used, overwritten).
template <int N>
__device__ __forceinline__ bool chol_solve(double (&A)[N][N], double (&b)[N])
{
    bool ok = true;
#pragma unroll
    for (int j = 0; j < N; ++j) {
        double d = A[j][j];
#pragma unroll
        for (int k = 0; k < j; ++k) d -= A[k][j] * A[k][j];
        ok = ok && (d > 0);
        d = sqrt(d);
        A[j][j] = d;
        const double id = 1.0 / d;
#pragma unroll
        for (int i = j + 1; i < N; ++i) {
            double v = A[j][i];
#pragma unroll
            for (int k = 0; k < j; ++k) v -= A[k][j] * A[k][i];
            A[j][i] = v * id;
        }
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
        double v = b[i];
#pragma unroll
        for (int k = 0; k < i; ++k) v -= A[k][i] * b[k];
        b[i] = v / A[i][i];
    }
#pragma unroll
    for (int i = N - 1; i >= 0; --i) {
        double v = b[i];
#pragma unroll
        for (int k = i + 1; k < N; ++k) v -= A[i][k] * b[k];
        b[i] = v / A[i][i];
    }
    return ok;
}

__device__ __forceinline__ void rodrigues_exp(const double* w, double* R)
{
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = sqrt(th2);
    double A, B;
    if (th < 1e-8) { A = 1.0 - th2 / 6.0; B = 0.5 - th2 / 24.0; }
    else { double s, c; sincos(th, &s, &c); A = s / th; B = (1.0 - c) / th2; }
    const double x = w[0], y = w[1], z = w[2];
    R[0] = 1.0 - B * (y * y + z * z); R[1] = B * x * y - A * z;         R[2] = B * x * z + A * y;
    R[3] = B * x * y + A * z;         R[4] = 1.0 - B * (x * x + z * z); R[5] = B * y * z - A * x;
    R[6] = B * x * z - A * y;         R[7] = B * y * z + A * x;         R[8] = 1.0 - B * (x * x + y * y);
}

// cv::Rodrigues' matrix -> vector branch structure (theta in [0, pi])
__device__ __forceinline__ void rodrigues_log(const double* R, double* w)
{
    double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
    const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
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
        theta /= sqrt(rx * rx + ry * ry + rz * rz);
        w[0] = rx * theta; w[1] = ry * theta; w[2] = rz * theta;
    } else {
        const double vth = theta / (2 * s);
        w[0] = rx * vth; w[1] = ry * vth; w[2] = rz * vth;
    }
}

struct Pose {
    double R[9], t[3];
};
struct Normal {  // J^T J (upper, row-major packed 21), J^T r (6), cost
    double a[21], g[6], cost;
};

struct Corr {
    const float* obj;  // m x 3
    const float* img;  // m x 2
    int m;
    double fx, fy, cx, cy;
};

// one pass over the correspondences at pose P; result identical on every lane
__device__ __forceinline__ void normal_equations(const Corr& c, const Pose& P, int lane, double* xch, Normal& out)
{
    double v[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) v[k] = 0;
    for (int i = lane; i < c.m; i += 32) {
        const double X = c.obj[3 * i], Y = c.obj[3 * i + 1], Z = c.obj[3 * i + 2];
        const double qx = P.R[0] * X + P.R[1] * Y + P.R[2] * Z, qy = P.R[3] * X + P.R[4] * Y + P.R[5] * Z,
                     qz = P.R[6] * X + P.R[7] * Y + P.R[8] * Z;
        const double px = qx + P.t[0], py = qy + P.t[1], pz = qz + P.t[2];
        const double iz = 1.0 / pz;
        const double ru = c.fx * px * iz + c.cx - (double)c.img[2 * i], rv = c.fy * py * iz + c.cy - (double)c.img[2 * i + 1];
        const double a0 = c.fx * iz, a2 = -c.fx * px * iz * iz, b1 = c.fy * iz, b2 = -c.fy * py * iz * iz;
        double Ju[6], Jv[6];
        Ju[0] = a2 * qy;            Ju[1] = a0 * qz - a2 * qx; Ju[2] = -a0 * qy; Ju[3] = a0; Ju[4] = 0;  Ju[5] = a2;
        Jv[0] = -b1 * qz + b2 * qy; Jv[1] = -b2 * qx;          Jv[2] = b1 * qx;  Jv[3] = 0;  Jv[4] = b1; Jv[5] = b2;
        int e = 0;
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int cc = r; cc < 6; ++cc) v[e++] += Ju[r] * Ju[cc] + Jv[r] * Jv[cc];
#pragma unroll
        for (int r = 0; r < 6; ++r) v[21 + r] += Ju[r] * ru + Jv[r] * rv;
        v[27] += ru * ru + rv * rv;
    }
    warp_transpose_sum32(v, lane);
    __syncwarp();
    xch[lane] = v[0];
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 21; ++k) out.a[k] = xch[k];
#pragma unroll
    for (int k = 0; k < 6; ++k) out.g[k] = xch[21 + k];
    out.cost = xch[27];
    __syncwarp();
}

// LM from P; returns passes.  Every lane runs the same scalar control flow on identical values.
__device__ int refine(const Corr& c, Pose& P, int max_iters, int lane, double* xch, double& cost_out)
{
    Normal N;
    normal_equations(c, P, lane, xch, N);
    double lambda = 1e-3;
    int passes = 1;
    for (int it = 0; it < max_iters; ++it) {
        double A[6][6], d[6];
        int e = 0;
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int cc = r; cc < 6; ++cc) { A[r][cc] = N.a[e]; A[cc][r] = N.a[e]; ++e; }
#pragma unroll
        for (int k = 0; k < 6; ++k) { A[k][k] += lambda * A[k][k]; d[k] = -N.g[k]; }
        if (!chol_solve<6>(A, d)) {
            lambda *= 10;
            if (lambda > 1e12) break;
            continue;
        }
        Pose Q;
        double dR[9];
        rodrigues_exp(d, dR);
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) Q.R[i * 3 + j] = dR[i * 3] * P.R[j] + dR[i * 3 + 1] * P.R[3 + j] + dR[i * 3 + 2] * P.R[6 + j];
#pragma unroll
        for (int k = 0; k < 3; ++k) Q.t[k] = P.t[k] + d[3 + k];
        Normal M;
        normal_equations(c, Q, lane, xch, M);
        ++passes;
        if (M.cost < N.cost) {
            double step = 0;
#pragma unroll
            for (int k = 0; k < 6; ++k) step = fmax(step, fabs(d[k]));
            const double gain = N.cost - M.cost;
            P = Q;
            N = M;
            lambda = fmax(lambda * 0.1, 1e-15);
            if (step < 1e-9 || gain <= 1e-12 * N.cost) break;
        } else {
            lambda *= 10;
            if (lambda > 1e12) break;
        }
    }
    cost_out = N.cost;
    return passes;
}

// homography initialisation for a planar (z = 0) object
__device__ bool init_planar(const Corr& c, Pose& P, int lane, double* xch)
{
    double sx = 0, sy = 0, sxx = 0, syy = 0;
    for (int i = lane; i < c.m; i += 32) {
        const double X = c.obj[3 * i], Y = c.obj[3 * i + 1];
        sx += X; sy += Y; sxx += X * X; syy += Y * Y;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o); sy += __shfl_xor_sync(0xffffffffu, sy, o);
        sxx += __shfl_xor_sync(0xffffffffu, sxx, o); syy += __shfl_xor_sync(0xffffffffu, syy, o);
    }
    const double Xc = sx / c.m, Yc = sy / c.m, var = (sxx / c.m - Xc * Xc) + (syy / c.m - Yc * Yc);
    if (!(var > 0)) return false;
    const double s = 1.0 / sqrt(var);
    // 36 + 8 sums in two transposing reductions
    double u[32], w[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) { u[k] = 0; w[k] = 0; }
    for (int i = lane; i < c.m; i += 32) {
        const double X = s * ((double)c.obj[3 * i] - Xc), Y = s * ((double)c.obj[3 * i + 1] - Yc);
        const double x = ((double)c.img[2 * i] - c.cx) / c.fx, y = ((double)c.img[2 * i + 1] - c.cy) / c.fy;
        const double a1[8] = {X, Y, 1, 0, 0, 0, -x * X, -x * Y}, a2[8] = {0, 0, 0, X, Y, 1, -y * X, -y * Y};
        int e = 0;
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int cc = r; cc < 8; ++cc) {
                const double p = a1[r] * a1[cc] + a2[r] * a2[cc];
                if (e < 32) u[e] += p; else w[e - 32] += p;
                ++e;
            }
#pragma unroll
        for (int r = 0; r < 8; ++r) w[4 + r] += a1[r] * x + a2[r] * y;
    }
    warp_transpose_sum32(u, lane);
    warp_transpose_sum32(w, lane);
    __syncwarp();
    xch[lane] = u[0];
    xch[32 + lane] = w[0];
    __syncwarp();
    double M[8][8], b[8];
    {
        int e = 0;
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int cc = r; cc < 8; ++cc) { M[r][cc] = xch[e]; M[cc][r] = xch[e]; ++e; }
#pragma unroll
        for (int r = 0; r < 8; ++r) b[r] = xch[36 + r];
    }
    __syncwarp();
    if (!chol_solve<8>(M, b)) return false;
    double h1[3] = {s * b[0], s * b[3], s * b[6]}, h2[3] = {s * b[1], s * b[4], s * b[7]};
    double h3[3] = {b[2] - Xc * h1[0] - Yc * h2[0], b[5] - Xc * h1[1] - Yc * h2[1], 1.0 - Xc * h1[2] - Yc * h2[2]};
    if (h3[2] < 0)
        for (int k = 0; k < 3; ++k) { h1[k] = -h1[k]; h2[k] = -h2[k]; h3[k] = -h3[k]; }
    const double n1 = sqrt(h1[0] * h1[0] + h1[1] * h1[1] + h1[2] * h1[2]), n2 = sqrt(h2[0] * h2[0] + h2[1] * h2[1] + h2[2] * h2[2]);
    if (!(n1 > 0 && n2 > 0)) return false;
    const double ts = 2.0 / (n1 + n2);
    double r1[3], r2[3], r3[3];
    for (int k = 0; k < 3; ++k) { r1[k] = h1[k] / n1; r2[k] = h2[k] / n2; P.t[k] = h3[k] * ts; }
    const double d = r1[0] * r2[0] + r1[1] * r2[1] + r1[2] * r2[2];
    for (int k = 0; k < 3; ++k) r2[k] -= d * r1[k];
    const double n = sqrt(r2[0] * r2[0] + r2[1] * r2[1] + r2[2] * r2[2]);
    if (!(n > 0)) return false;
    for (int k = 0; k < 3; ++k) r2[k] /= n;
    r3[0] = r1[1] * r2[2] - r1[2] * r2[1]; r3[1] = r1[2] * r2[0] - r1[0] * r2[2]; r3[2] = r1[0] * r2[1] - r1[1] * r2[0];
    for (int k = 0; k < 3; ++k) { P.R[k * 3] = r1[k]; P.R[k * 3 + 1] = r2[k]; P.R[k * 3 + 2] = r3[k]; }
    return true;
}

struct PnpArgs {
    const float* obj;  // n x m x 3 (or m x 3 shared by all frames when obj_stride == 0)
    const float* img;  // n x m x 2
    size_t obj_stride, img_stride;  // floats between frames
    int n, m;
    const float* K;     // n x 9 or 9 shared (k_stride == 0)
    int k_stride;
    const double* guess;  // n x 6 or null
    int flags, max_iters;
    mantis_pnp_result* out;
};

__global__ void __launch_bounds__(PNT) k_pnp(PnpArgs a)
{
    __shared__ double s_xch[PNT / 32][64];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int frame = blockIdx.x * (PNT / 32) + warp;
    if (frame >= a.n) return;
    double* xch = s_xch[warp];
    Corr c;
    c.obj = a.obj + (size_t)frame * a.obj_stride;
    c.img = a.img + (size_t)frame * a.img_stride;
    c.m = a.m;
    const float* K = a.K + (size_t)frame * a.k_stride;
    c.fx = K[0]; c.fy = K[4]; c.cx = K[2]; c.cy = K[5];
    mantis_pnp_result* o = a.out + frame;

    Pose P;
    bool ok = a.m >= 4;
    if (ok) {
        if (a.guess) {
            const double* g = a.guess + (size_t)frame * 6;
            rodrigues_exp(g, P.R);
            P.t[0] = g[3]; P.t[1] = g[4]; P.t[2] = g[5];
        } else {
            ok = init_planar(c, P, lane, xch);
        }
    }
    double cost = 0;
    int passes = 0;
    if (ok) {
        passes = refine(c, P, a.max_iters, lane, xch, cost);
        if (a.flags & MANTIS_PNP_BOTH_PLANAR) {
            // second planar candidate: the plane normal mirrored about the line of sight to the centroid
            double cen[3] = {0, 0, 0}, oc[3] = {0, 0, 0};
            for (int i = lane; i < c.m; i += 32)
                for (int k = 0; k < 3; ++k) oc[k] += c.obj[3 * i + k];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
#pragma unroll
                for (int of = 16; of > 0; of >>= 1) oc[k] += __shfl_xor_sync(0xffffffffu, oc[k], of);
                oc[k] /= c.m;
            }
            for (int k = 0; k < 3; ++k) cen[k] = P.R[k * 3] * oc[0] + P.R[k * 3 + 1] * oc[1] + P.R[k * 3 + 2] * oc[2] + P.t[k];
            const double cn = sqrt(cen[0] * cen[0] + cen[1] * cen[1] + cen[2] * cen[2]);
            const double cd[3] = {cen[0] / cn, cen[1] / cn, cen[2] / cn}, nn[3] = {P.R[2], P.R[5], P.R[8]};
            const double ax[3] = {nn[1] * cd[2] - nn[2] * cd[1], nn[2] * cd[0] - nn[0] * cd[2], nn[0] * cd[1] - nn[1] * cd[0]};
            const double sn = sqrt(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]), cs = nn[0] * cd[0] + nn[1] * cd[1] + nn[2] * cd[2];
            if (sn > 1e-6) {
                const double ang = 2.0 * atan2(sn, cs);
                const double w[3] = {ax[0] / sn * ang, ax[1] / sn * ang, ax[2] / sn * ang};
                double dR[9];
                rodrigues_exp(w, dR);
                Pose Q;
                for (int i = 0; i < 3; ++i)
                    for (int j = 0; j < 3; ++j) Q.R[i * 3 + j] = dR[i * 3] * P.R[j] + dR[i * 3 + 1] * P.R[3 + j] + dR[i * 3 + 2] * P.R[6 + j];
                for (int k = 0; k < 3; ++k) Q.t[k] = cen[k] - (Q.R[k * 3] * oc[0] + Q.R[k * 3 + 1] * oc[1] + Q.R[k * 3 + 2] * oc[2]);
                double cost2;
                passes += refine(c, Q, a.max_iters, lane, xch, cost2);
                if (cost2 < cost) { P = Q; cost = cost2; }
            }
        }
    }
    if (lane == 0) {
        if (ok) {
            rodrigues_log(P.R, o->rvec);
            o->tvec[0] = P.t[0]; o->tvec[1] = P.t[1]; o->tvec[2] = P.t[2];
            o->reproj_error = sqrt(cost);
        } else {
            for (int k = 0; k < 3; ++k) { o->rvec[k] = 0; o->tvec[k] = 0; }
            o->reproj_error = 0;
        }
        o->passes = passes;
        o->status = ok ? MANTIS_PNP_OK : MANTIS_PNP_DEGENERATE;
    }
}

}  // namespace

extern "C" int mantis_pnp_batch(mantis_ctx* c, const float* obj_xyz, const float* img_xy, int n, int m, int obj_shared,
                                const float* K, int k_shared, const double* guess_rt, const mantis_pnp_params* prm,
                                mantis_pnp_result* out, int mem)
{
    if (!c || n < 0 || m < 0 || (n > 0 && (!obj_xyz || !img_xy || !K || !out)) ||
        (mem != MANTIS_MEM_HOST && mem != MANTIS_MEM_DEVICE))
        return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    PnpArgs a;
    a.n = n; a.m = m;
    a.obj_stride = obj_shared ? 0 : (size_t)3 * m;
    a.img_stride = (size_t)2 * m;
    a.k_stride = k_shared ? 0 : 9;
    a.flags = prm ? prm->flags : 0;
    a.max_iters = prm && prm->max_iters > 0 ? prm->max_iters : 50;
    a.obj = obj_xyz; a.img = img_xy; a.K = K; a.guess = guess_rt; a.out = out;
    if (mem == MANTIS_MEM_HOST) {
        const size_t b_obj = sizeof(float) * 3 * m * (obj_shared ? 1 : (size_t)n), b_img = sizeof(float) * 2 * m * (size_t)n;
        const size_t b_k = sizeof(float) * 9 * (k_shared ? 1 : (size_t)n), b_g = guess_rt ? sizeof(double) * 6 * (size_t)n : 0;
        const size_t b_out = sizeof(mantis_pnp_result) * (size_t)n;
        auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
        const size_t total = up(b_obj) + up(b_img) + up(b_k) + up(b_g) + up(b_out);
        int rc;
        if ((rc = ctx_reserve(c, c->pnp_stage, total))) return rc;
        char* p = (char*)c->pnp_stage.p;
        MCHECK(c, cudaMemcpyAsync(p, obj_xyz, b_obj, cudaMemcpyHostToDevice, c->stream)); a.obj = (const float*)p; p += up(b_obj);
        MCHECK(c, cudaMemcpyAsync(p, img_xy, b_img, cudaMemcpyHostToDevice, c->stream)); a.img = (const float*)p; p += up(b_img);
        MCHECK(c, cudaMemcpyAsync(p, K, b_k, cudaMemcpyHostToDevice, c->stream)); a.K = (const float*)p; p += up(b_k);
        if (guess_rt) { MCHECK(c, cudaMemcpyAsync(p, guess_rt, b_g, cudaMemcpyHostToDevice, c->stream)); a.guess = (const double*)p; }
        p += up(b_g);
        a.out = (mantis_pnp_result*)p;
    }
    k_pnp<<<(n + PNT / 32 - 1) / (PNT / 32), PNT, 0, c->stream>>>(a);
    c->launches++;
    MCHECK(c, cudaGetLastError());
    if (mem == MANTIS_MEM_HOST)
        MCHECK(c, cudaMemcpyAsync(out, a.out, sizeof(mantis_pnp_result) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}
