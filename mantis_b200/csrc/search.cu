// search.cu -- K8/K9/K10: map-correspondence scoring, stochastic refinement, cell shifts and yaw
// disambiguation; one 1024-thread CTA per frame.
//
//   evaluateHypotheses / evaluateHypothesis / evaluateHypothesisWithImageWHITE / computePointErrorFAST
//                                     include/mantis3/HypothesisEvaluation.h:31-41,71-158,218-227
//   getBestNHypotheses                :484-518
//   optimizeHypothesisWithParticleFilter, generateRandomHypothesis   PoseAdjustment.h:13-60
//   computeAllShiftedHypothesesFAST, shiftHypothesis                 HypothesisGeneration.h:111-140
//   determineBestYaw, evaluateHypothesesColor, evaluateHypothesisWithImageCOLOR, computePointError
//                                     HypothesisEvaluation.h:43-66,89-105,160-266,521-581
//
// Scoring is a warp task: 32 map points of one hypothesis are projected in fp64 through the
// fisheye model, the pixel (cvRound addressing, mask bit) is gathered from the L2-resident frame,
// and the integer colour error is shuffled down to one shared-memory atomic per warp.  Integer
// sums make the score independent of evaluation order, so the greedy `<` decisions are exact.
// The Gaussian perturbations of the refinement are a fixed table: the reference draws them from
// cv::RNG(1), here re-seeded per frame, so they are the same for every frame of a context.
#include "ctx.h"
#include "posemath.cuh"

namespace {

constexpr int SNT = 1024;
constexpr int NWARP = SNT / 32;
constexpr int MAX_MAP = 1024;
constexpr double DBLMAX = 1.7976931348623157e308;

struct Img {
    const uint8_t* bgr;
    const uint32_t* mask;  // 1 bit per pixel, row-major words
    int W, H, WW, stride;
};

// img.at<Vec3b>(Point(x, y)) with the reference's unchecked linear addressing over a tightly
// packed image: pixel index y*W + x; anything outside the buffer reads black.
__device__ __forceinline__ int pixel_err(const Img& im, int x, int y, bool use_mask, int d0, int d1, int d2)
{
    long long pix = (long long)y * im.W + x;
    int b = 0, g = 0, r = 0;
    if (pix >= 0 && pix < (long long)im.W * im.H) {
        int yy = y, xx = x;
        if ((unsigned)x >= (unsigned)im.W) { yy = (int)(pix / im.W); xx = (int)(pix - (long long)yy * im.W); }
        bool keep = true;
        if (use_mask) keep = (__ldg(im.mask + (size_t)yy * im.WW + (xx >> 5)) >> (xx & 31)) & 1u;
        if (keep) {
            const uint8_t* p = im.bgr + (size_t)yy * im.stride + 3 * xx;
            b = __ldg(p); g = __ldg(p + 1); r = __ldg(p + 2);
        }
    }
    int db = b - d0, dg = g - d1, dr = r - d2;
    return db * db + dg * dg + dr * dr;
}

// Hypothesis::projectPoint + distortPixel + inFrame
__device__ __forceinline__ bool project(const Tf& c2w, const pm::Camera& cam, const double* p, int W, int H, double& u,
                                        double& v)
{
    double c[3];
    for (int i = 0; i < 3; ++i) c[i] = (c2w.R[i * 3] * p[0] + c2w.R[i * 3 + 1] * p[1] + c2w.R[i * 3 + 2] * p[2]) + c2w.t[i];
    if (!(c[2] > 0)) return false;
    pm::fisheye_distort(cam, c[0] / c[2], c[1] / c[2], u, v);
    return (u < W && v < H && u >= 0 && v > 0);
}

struct SearchArgs {
    const uint8_t* bgr; size_t pitch; int stride;
    const uint32_t* mask_bits;
    int W, H, WW;
    const pm::Camera* cams;
    const FrameQuads* quads;
    const Hyp* hyps; const short* sel; const FramePose* info;
    const double* map;  // xyz, white | red | green
    int n_map, n_green_off, n_green;
    const Tf* rand_tf;  // [iterations * particles]
    const double* shifts;  // [n_shift][2]
    int n_shift, iterations, particles;
    Tf rotz;
    double projection_bias, color_bias;
    mantis_result* out;
    mantis_frame_trace* trace;  // may be null
};

struct Shared {
    double map[MAX_MAP * 3];
    unsigned long long sum[1024];
    int cnt[1024];
    double err[1024];
    Tf cand_w2c[96], cand_c2w[96];
    Hyp best;
    double green_val[80 * 40];
    unsigned char green_ok[80 * 40];
    double hyp_err[96];
    short top20[20];
};

// score `nh` hypotheses (c2w via the accessor) against all map points on the masked image
template <class GetC2W>
__device__ void score_white(Shared& s, const Img& im, const pm::Camera& cam, int n_map, int nh, double bias, GetC2W get)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < nh; i += SNT) { s.sum[i] = 0ull; s.cnt[i] = 0; }
    __syncthreads();
    const int chunks = (n_map + 31) >> 5;
    for (int task = warp; task < nh * chunks; task += NWARP) {
        const int h = task / chunks, ch = task - h * chunks;
        const Tf& c2w = get(h);
        const int pi = ch * 32 + lane;
        int e = 0, ok = 0;
        if (pi < n_map) {
            double u, v;
            if (project(c2w, cam, &s.map[3 * pi], im.W, im.H, u, v)) {
                ok = 1;
                e = pixel_err(im, __double2int_rn(u), __double2int_rn(v), true, 255, 255, 255);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            e += __shfl_xor_sync(0xffffffffu, e, o);
            ok += __shfl_xor_sync(0xffffffffu, ok, o);
        }
        if (lane == 0 && ok) {
            atomicAdd(&s.sum[h], (unsigned long long)e);
            atomicAdd(&s.cnt[h], ok);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nh; i += SNT)
        s.err[i] = s.cnt[i] > 0 ? (double)s.sum[i] / ((double)s.cnt[i] * bias) : DBLMAX;
    __syncthreads();
}

__global__ void __launch_bounds__(SNT) k_search(SearchArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Shared& s = *reinterpret_cast<Shared*>(smem_raw);
    const int frame = blockIdx.x, tid = threadIdx.x;
    mantis_result* out = a.out + frame;
    mantis_frame_trace* tr = a.trace ? a.trace + frame : nullptr;
    const FrameQuads* fq = a.quads + frame;
    const FramePose* info = a.info + frame;
    const int n_sel = info->n_sel;

    if (tid == 0) {
        out->n_quads = fq->n_quads;
        out->n_hyps = n_sel;
        out->n_particles = 0;
        out->error = 0; out->min_yaw_diff = 0;
        for (int k = 0; k < 3; ++k) out->pos[k] = 0;
        for (int k = 0; k < 4; ++k) out->quat_xyzw[k] = 0;
        out->status = fq->flags ? MANTIS_FRAME_OVERFLOW
                    : fq->n_quads == 0 ? MANTIS_FRAME_NO_QUADS
                    : n_sel == 0 ? MANTIS_FRAME_NO_HYPOTHESES : MANTIS_FRAME_OK;
        if (tr) {
            tr->n_contours = fq->n_contours; tr->n_quads_raw = fq->n_quads_raw; tr->n_quads = fq->n_quads;
            tr->flags = fq->flags; tr->n_hyps_generated = info->n_hyps_generated; tr->n_hyps_cluster = n_sel;
            tr->best_rot = -1; tr->pad_ = 0;
        }
    }
    if (fq->flags || fq->n_quads == 0 || n_sel == 0) return;

    Img im;
    im.bgr = a.bgr + (size_t)frame * a.pitch;
    im.mask = a.mask_bits + (size_t)frame * a.H * a.WW;
    im.W = a.W; im.H = a.H; im.WW = a.WW; im.stride = a.stride;
    const pm::Camera cam = a.cams[frame];
    const Hyp* hyps = a.hyps + (size_t)frame * MANTIS_MAX_HYPS;
    const short* sel = a.sel + (size_t)frame * MANTIS_MAX_HYPS;
    for (int i = tid; i < a.n_map * 3; i += SNT) s.map[i] = a.map[i];
    __syncthreads();

    // (6) evaluateHypotheses on the kept cluster, (7) best 1
    score_white(s, im, cam, a.n_map, n_sel, a.projection_bias, [&](int h) -> const Tf& { return hyps[sel[h]].c2w; });
    if (tid == 0) {
        int bi = 0;  // stable descending sort, last element: minimal error, highest index among ties
        for (int i = 1; i < n_sel; ++i)
            if (s.err[i] <= s.err[bi]) bi = i;
        s.best = hyps[sel[bi]];
        s.best.error = s.err[bi];
        if (tr) tr->first_best_error = s.err[bi];
    }
    __syncthreads();

    // (8) optimizeHypothesisWithParticleFilter
    for (int it = 0; it < a.iterations; ++it) {
        if (tid < a.particles) {
            pm::tf_mul(s.best.w2c, a.rand_tf[it * a.particles + tid], s.cand_w2c[tid]);
            pm::tf_inv(s.cand_w2c[tid], s.cand_c2w[tid]);
        }
        __syncthreads();
        score_white(s, im, cam, a.n_map, a.particles, a.projection_bias, [&](int h) -> const Tf& { return s.cand_c2w[h]; });
        if (tid == 0) {
            int bp = -1;
            double be = s.best.error;
            for (int p = 0; p < a.particles; ++p)
                if (s.err[p] < be) { be = s.err[p]; bp = p; }
            if (bp >= 0) {
                s.best.w2c = s.cand_w2c[bp];
                s.best.c2w = s.cand_c2w[bp];
                s.best.error = be;
            }
        }
        __syncthreads();
    }
    if (tid == 0 && tr) {
        tr->optimized_error = s.best.error;
        for (int k = 0; k < 9; ++k) tr->optimized_w2c[k] = s.best.w2c.R[k];
        for (int k = 0; k < 3; ++k) tr->optimized_w2c[9 + k] = s.best.w2c.t[k];
    }

    // (9) the 81 cell shifts, (10) evaluate
    const int ns = a.n_shift;
    if (tid < ns) {
        Tf w = s.best.w2c;
        w.t[0] += a.shifts[2 * tid];
        w.t[1] += a.shifts[2 * tid + 1];
        w.t[2] += 0;
        s.cand_w2c[tid] = w;
        pm::tf_inv(w, s.cand_c2w[tid]);
    }
    __syncthreads();
    score_white(s, im, cam, a.n_map, ns, a.projection_bias, [&](int h) -> const Tf& { return s.cand_c2w[h]; });
    if (tr)
        for (int i = tid; i < ns && i < 81; i += SNT) tr->shift_err[i] = s.err[i];
    // (11) best 20: position in a stable descending sort; keep the last 20
    const int keep = ns < 20 ? ns : 20;
    if (tid < ns) {
        int pos = 0;
        const double e = s.err[tid];
        for (int j = 0; j < ns; ++j) pos += (s.err[j] > e) || (s.err[j] == e && j < tid);
        if (pos >= ns - keep) s.top20[pos - (ns - keep)] = (short)tid;
    }
    __syncthreads();

    // (12) determineBestYaw on the original image: 4 rotations x `keep` hypotheses x green points
    // rotated copies: slot r*keep + k; reuse cand arrays beyond the shifts' slots is unsafe, so
    // stage the kept shifts first
    __shared__ Tf s_rot_w2c[80], s_rot_c2w[80];
    if (tid < keep) {
        Tf w = s.cand_w2c[s.top20[tid]];
        for (int r = 0; r < 4; ++r) {
            if (r > 0) { Tf t2; pm::tf_mul(a.rotz, w, t2); w = t2; }
            s_rot_w2c[r * keep + tid] = w;
            pm::tf_inv(w, s_rot_c2w[r * keep + tid]);
        }
    }
    __syncthreads();
    const int ng = a.n_green, nrot = 4 * keep;
    for (int item = tid; item < nrot * ng; item += SNT) {
        const int h = item / ng, gi = item - h * ng;
        double u, v;
        bool ok = project(s_rot_c2w[h], cam, &s.map[3 * (a.n_green_off + gi)], im.W, im.H, u, v);
        double e = 0;
        if (ok) {
            long long acc = 0;
            for (int dx = -5; dx < 5; ++dx)
                for (int dy = -5; dy < 5; ++dy)
                    acc += pixel_err(im, __double2int_rn(u + (double)dx), __double2int_rn(v + (double)dy), false, 50, 255, 85);
            e = (double)acc / 100.0;
        }
        s.green_val[h * 40 + gi] = e;
        s.green_ok[h * 40 + gi] = ok;
    }
    __syncthreads();
    if (tid < nrot) {
        double err = 0;
        int n = 0;
        for (int gi = 0; gi < ng; ++gi)
            if (s.green_ok[tid * 40 + gi]) { err += s.green_val[tid * 40 + gi]; ++n; }
        s.hyp_err[tid] = n > 0 ? err / ((double)n * a.color_bias) : DBLMAX;
    }
    __syncthreads();
    if (tid == 0) {
        double yaw[4], best_error = DBLMAX;
        int best = -1;
        for (int r = 0; r < 4; ++r) {
            double total = 0;
            int success = 0;
            for (int k = 0; k < keep; ++k) {
                double e = s.hyp_err[r * keep + k];
                if (fabs(e - DBLMAX) > 0.001) { success++; total += e; }
            }
            yaw[r] = success ? total / (double)success : DBLMAX;
            if (yaw[r] < best_error) { best_error = yaw[r]; best = r; }
        }
        double min1 = DBLMAX, min2 = DBLMAX;
        for (int r = 0; r < 4; ++r) {
            double diff = yaw[r] - best_error;
            if (diff < min1) { min2 = min1; min1 = diff; }
            else if (diff < min2) min2 = diff;
        }
        out->min_yaw_diff = min2;
        out->n_particles = n_sel + 1 + a.iterations * a.particles + ns + nrot;
        if (tr) { for (int r = 0; r < 4; ++r) tr->yaw_err[r] = yaw[r]; tr->best_rot = best; }
        if (best < 0) {
            out->status = MANTIS_FRAME_YAW_FAILED;
        } else {
            const Tf& w = s_rot_w2c[best * keep + keep - 1];  // hyps.back()
            for (int k = 0; k < 3; ++k) out->pos[k] = w.t[k];
            pm::mat_to_quat(w.R, out->quat_xyzw);
            out->error = s.hyp_err[best * keep + keep - 1];
        }
    }
}

}  // namespace

int launch_search(mantis_ctx* c, const uint8_t* d_bgr, size_t pitch, int stride, const uint32_t* d_mask_bits, int n,
                  int w, int h, const pm::Camera* d_cams, const FrameQuads* d_quads, const Hyp* d_hyps,
                  const short* d_sel, const FramePose* d_info, mantis_result* d_out, mantis_frame_trace* d_trace)
{
    if (c->n_map <= 0 || c->n_map > MAX_MAP || c->n_green > 40 || c->n_shift > 96 || c->cfg.particles > 96) {
        snprintf(c->err, sizeof(c->err), "map / search sizes exceed the kernel's limits (map %d, green %d)", c->n_map, c->n_green);
        return MANTIS_ERR_INVALID;
    }
    SearchArgs a;
    a.bgr = d_bgr; a.pitch = pitch; a.stride = stride;
    a.mask_bits = d_mask_bits;
    a.W = w; a.H = h; a.WW = words_per_row(w);
    a.cams = d_cams; a.quads = d_quads; a.hyps = d_hyps; a.sel = d_sel; a.info = d_info;
    a.map = (const double*)c->d_map.p;
    a.n_map = c->n_map; a.n_green_off = c->n_map - c->n_green; a.n_green = c->n_green;
    a.rand_tf = (const Tf*)c->d_rand.p;
    a.shifts = (const double*)c->d_shifts.p;
    a.n_shift = c->n_shift; a.iterations = c->cfg.iterations; a.particles = c->cfg.particles;
    a.rotz = c->rotz;
    a.projection_bias = c->cfg.projection_bias;
    a.color_bias = c->cfg.color_projection_bias;
    a.out = d_out; a.trace = d_trace;
    const size_t smem = sizeof(Shared);
    MCHECK(c, cudaFuncSetAttribute(k_search, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_search<<<n, SNT, smem, c->stream>>>(a);
    c->launches++;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}
