// search.cu -- K8/K9/K10: map-correspondence scoring, stochastic refinement, cell shifts and yaw
// disambiguation; persistent 1024-thread CTAs (one per SM), frames handed out through a ticket counter.
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
// fisheye model (branch-free), the pixel (cvRound addressing; mask bit from the frame's mask plane,
// staged in shared memory when it fits) is gathered from the L2-resident frame, and the integer colour
// error is reduced to one shared-memory atomic per warp and hypothesis.  Tasks run in blocks of four:
// projections first, then the block's mask words, then its pixel words (score_loop).  Integer sums
// make the score independent of evaluation order, so the greedy `<` decisions are exact.  Every scoring
// phase of a frame goes through one inlined copy of that loop (phase loop in k_search).
// The Gaussian perturbations of the refinement are a fixed table: the reference draws them from
// cv::RNG(1), here re-seeded per frame, so they are the same for every frame of a context.
#include <cstddef>
#include <cstdlib>
#include "ctx.h"
#include "fastmath.cuh"
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

constexpr int ATAN_N = 128;  // table step h = pi/256 over [0, pi/2)
// atan(r) for r >= 0 with inv_r = 1/r.  The table index only has to land within one step of the
// angle, so it comes from a 3-term fp32 estimate (|error| < 0.004 rad = 0.31 h).
__device__ __forceinline__ double fast_atan_pos(double r, double inv_r, const double2* __restrict__ tab)
{
    const bool big = r > 1.0;
    const float q = (float)(big ? inv_r : r);
    float est = q * fmaf(0.273f, 1.0f - q, 0.78539816f);
    if (big) est = 1.57079633f - est;
    const int i = min(__float2int_rn(est * 81.48733086305041f), ATAN_N - 1);  // 256/pi
    const double2 e = tab[i];
    const double x = (r - e.y) * fast_rcp3(fma(r, e.y, 1.0));
    const double x2 = x * x;
    const double p = fma(x2, fma(x2, fma(x2, -1.0 / 7.0, 0.2), -1.0 / 3.0), 1.0);
    return fma(x, p, e.x);
}

// img.at<Vec3b>(Point(x, y)) with the reference's unchecked linear addressing over a tightly
// packed image: pixel index y*W + x; anything outside the buffer reads black.
__device__ __forceinline__ int pixel_err(const Img& im, int x, int y, bool use_mask, int d0, int d1, int d2)
{
    int b = 0, g = 0, r = 0;
    if ((unsigned)x < (unsigned)im.W && (unsigned)y < (unsigned)im.H) {  // the common case
        bool keep = true;
        if (use_mask) keep = (__ldg(im.mask + (unsigned)(y * im.WW + (x >> 5))) >> (x & 31)) & 1u;
        if (keep) {
            const uint8_t* p = im.bgr + (unsigned)(y * im.stride + 3 * x);
            b = __ldg(p); g = __ldg(p + 1); r = __ldg(p + 2);
        }
        int db = b - d0, dg = g - d1, dr = r - d2;
        return db * db + dg * dg + dr * dr;
    }
    long long pix = (long long)y * im.W + x;
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

// Hypothesis::projectPoint + distortPixel (Kannala-Brandt, Mantis3Types.h:88-136) + inFrame
// (HypothesisEvaluation.h:388-398).  Same formulae as pm::fisheye_distort, with the fast fp64 blocks.
struct CamK {
    double fx, fy, cx, cy, D0, D1, D2, D3;
    bool has_dist;
};
struct PoseXY {  // the part of a camera<-world transform that a floor-plane point needs, plus where the rest lives
    double rx[3], ry[3], t[3];
    const Tf* full;
};
__device__ __forceinline__ PoseXY load_pose(const Tf& c2w)
{
    PoseXY p;
#pragma unroll
    for (int i = 0; i < 3; ++i) { p.rx[i] = c2w.R[i * 3]; p.ry[i] = c2w.R[i * 3 + 1]; p.t[i] = c2w.t[i]; }
    p.full = &c2w;
    return p;
}
__device__ __forceinline__ bool project(const PoseXY& P, const CamK& cam, double px, double py, double pz, bool planar,
                                        const double2* __restrict__ atan_tab, int W, int H, double& u, double& v)
{
    double c[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double acc = fma(P.rx[i], px, fma(P.ry[i], py, P.t[i]));
        if (!planar) acc = fma(P.full->R[i * 3 + 2], pz, acc);  // the map is the floor plane: z = 0 adds nothing
        c[i] = acc;
    }
    if (!(c[2] > 1e-300)) return false;  // behind the camera (the reference's z > 0), or no usable reciprocal
    const double iz = fast_rcp3(c[2]);
    const double x = c[0] * iz, y = c[1] * iz;
    const double r2 = fma(x, x, y * y);
    double cdist = 1;
    if (r2 > 1e-16 && r2 < 1e300) {  // r > 1e-8
        const double inv_r = fast_rsqrt3(r2), r = r2 * inv_r;
        const double theta = fast_atan_pos(r, inv_r, atan_tab);
        double theta_d = theta;
        if (cam.has_dist) {
            const double t2 = theta * theta;
            theta_d = theta * fma(t2, fma(t2, fma(t2, fma(t2, cam.D3, cam.D2), cam.D1), cam.D0), 1.0);
        }
        cdist = theta_d * inv_r;
    } else if (!(r2 <= 1e-16)) {
        return false;  // overflow / NaN: the reference's atan(inf)/inf gives NaN pixel coordinates, never in frame
    }
    u = fma(x * cdist, cam.fx, cam.cx);
    v = fma(y * cdist, cam.fy, cam.cy);
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
    int n_map, n_green_off, n_green, map_planar;
    const double2* atan_tab;  // (i pi/256, tan(i pi/256)), i < 128
    const Tf* rand_tf;  // [iterations * particles]
    const double* shifts;  // [n_shift][2]
    int n_shift, iterations, particles;
    // the shift hypotheses through the map's lattice (localize.cu: build_shift_tables): the distinct points p - shift,
    // and for (shift, map point) which of them it is.  n_shift_pts == 0: every shift is scored like any other hypothesis
    const double* shift_pts;          // [n_shift_pts][3]
    const unsigned short* shift_idx;  // [n_shift][chunks * 32], padded with n_shift_pts
    int n_shift_pts;
    double Wd, Hd;
    unsigned* ticket; int n;
    int mask_in_smem;  // the frame's mask plane fits behind `Shared` in dynamic shared memory
    Tf rotz;
    double projection_bias, color_bias;
    mantis_result* out;
    mantis_frame_trace* trace;  // may be null
    long long* dbg;             // tuning aid: clock64 per phase (null = off)
    // score-only mode (mantis_score_batch): ext_n caller-supplied poses per frame instead of the search
    // rigs (k_search_rig): a unit is C consecutive frames; hypotheses are base poses, camera k sees W2B * b2c[frame]
    int C;
    const Tf* b2c;              // per frame: the camera in the base frame (mantis_frame.cam_to_base)
    const double* ext_w2c;      // [n][ext_n][12] R row-major + t, what Hypothesis::setW2C receives; null = search
    double* ext_scores;         // [n][ext_n]
    int ext_n;
};

// What the scoring loop needs about the frame, in shared memory: every use is one broadcast LDS, which is
// cheaper than recomputing it from blockIdx or holding it in one of 64 registers.
struct ScoreFrame {
    double2 fcx, fcy;         // (fx, cx), (fy, cy)
    double D0, D1, D2, D3;
};

struct Gather {  // per frame: where the masked image lives
    const uint8_t* bgr_al;    // frame base rounded down to 4 bytes
    const uint32_t* mask;     // generic pointer: the copy of the frame's mask plane in shared memory when it fits, else global
    unsigned delta;           // frame base - bgr_al
};
struct Shared {
    double map_x[MAX_MAP], map_y[MAX_MAP], map_z[MAX_MAP];
    double2 atan_tab[ATAN_N];
    unsigned long long sum[1024];
    int cnt[1024];
    double err[1024];
    Tf cand_w2c[96], cand_c2w[96];
    Hyp best;
    Img img;  // per-frame image descriptor: kept here so that the base pointers are one broadcast load away
    ScoreFrame sf;
    Gather gat;
    int frame;
    double2 wpose[NWARP][8];  // per-warp staging of the current hypothesis: (rx,ry) x3, (t0,t1), (t2,-), (rz0,rz1), (rz2,-)
    unsigned int green_acc[80 * 40];  // integer colour error summed over the 10x10 window
    int green_xy[80 * 40];            // cvRound(u) | cvRound(v) << 16 of the projected point, -1 if not in frame
    unsigned char green_half[80 * 40];  // u or v exactly on a .5: per-pixel rounding differs from a shifted window
    double hyp_err[96];
    short top20[20];
};

// fp64 constants of the projection whose low word is not zero: as literals ptxas rebuilds each of them with
// two UMOVs at every use (12 issue slots per point); from the constant bank they are direct operands.
__constant__ double KC[6] = {1e-300, 1e-16, 1e300, -1.0 / 7.0, 0.2, -1.0 / 3.0};

// Branch-free projection (same operations, in the same order, as project()): every lane runs the whole
// chain and carries a validity predicate, so the projections of a block form straight-line code.  Garbage in (a point behind the camera) stays garbage and ends with ok = false.
template <bool PLANAR, bool DIST>
__device__ __forceinline__ bool project_bf(const double2* __restrict__ wp, const ScoreFrame& cam, double px, double py, double pz,
                                           const double2* __restrict__ tab, double Wd, double Hd, double& u, double& v)
{
    double c[3];
    {
        const double2 r0 = wp[0], r1 = wp[1], r2_ = wp[2], t01 = wp[3], t2 = wp[4];
        c[0] = fma(r0.x, px, fma(r0.y, py, t01.x));
        c[1] = fma(r1.x, px, fma(r1.y, py, t01.y));
        c[2] = fma(r2_.x, px, fma(r2_.y, py, t2.x));
        if (!PLANAR) {
            const double2 z01 = wp[5], z2 = wp[6];
            c[0] = fma(z01.x, pz, c[0]); c[1] = fma(z01.y, pz, c[1]); c[2] = fma(z2.x, pz, c[2]);
        }
    }
    bool ok = c[2] > KC[0];
    const double iz = fast_rcp3(c[2]);
    const double x = c[0] * iz, y = c[1] * iz;
    const double r2 = fma(x, x, y * y);
    const bool mid = r2 > KC[1];  // the reference's r > 1e-8; (r2 = 0 gives NaN below, the select still takes 1)
    ok = ok && r2 < KC[2];        // overflow; also drops NaN
    const double inv_r = fast_rsqrt3(r2), r = r2 * inv_r;
    // fast_atan_pos with an index that is safe for garbage input
    const bool big = r > 1.0;
    const float q = (float)(big ? inv_r : r);
    float est = q * fmaf(0.273f, 1.0f - q, 0.78539816f);
    if (big) est = 1.57079633f - est;
    const unsigned ti = min((unsigned)__float2int_rn(est * 81.48733086305041f), (unsigned)(ATAN_N - 1));
    const double2 e = tab[ti];
    const double ax = (r - e.y) * fast_rcp3(fma(r, e.y, 1.0));
    const double ax2 = ax * ax;
    const double ap = fma(ax2, fma(ax2, fma(ax2, KC[3], KC[4]), KC[5]), 1.0);
    const double theta = fma(ax, ap, e.x);
    double theta_d = theta;
    if (DIST) {
        const double t2 = theta * theta;
        theta_d = theta * fma(t2, fma(t2, fma(t2, fma(t2, cam.D3, cam.D2), cam.D1), cam.D0), 1.0);
    }
    const double cdist = mid ? theta_d * inv_r : 1.0;
    const double2 fcx = cam.fcx, fcy = cam.fcy;  // late broadcast loads instead of eight live registers
    u = fma(x * cdist, fcx.x, fcx.y);
    v = fma(y * cdist, fcy.x, fcy.y);
    return ok && u < Wd && v < Hd && u >= 0 && v > 0;
}

// x == W or y == H (the reference addresses pixel y*W + x unchecked): a few points per frame, kept out of line
__device__ __noinline__ void rare_point(Shared& s, int h, int x, int y)
{
    const int er = pixel_err(s.img, x, y, true, 255, 255, 255);
    if (er) atomicAdd(&s.sum[h], (unsigned long long)er);
}

// Score `nh` hypotheses against all map points on the masked image.  The c2w of hypothesis h is the Tf at
// tf_base + (sel ? sel[h] : h) * tf_stride (global or shared memory).
//
// A warp task is (hypothesis, 32 map points), projected in fp64, branch-free.  Tasks are cut into one contiguous
// range per warp and run in blocks of up to BLK tasks of one hypothesis: first the projections, then the mask
// words of the whole block, then its pixel words, then the colour errors.  The gathers of a block are
// independent of each other, so their latencies overlap without a software pipeline carried in registers.
// err = sum (255 - c)^2 is one IDP.4A on the complemented pixel.
template <bool PLANAR, bool DIST>
__device__ __forceinline__ void score_loop(Shared& s, const SearchArgs& a, int nh, const char* tf_base, int tf_stride,
                                           const short* sel)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_map = a.n_map;
    const int chunks = (n_map + 31) >> 5;
    const int total = nh * chunks, per = (total + NWARP - 1) / NWARP;
    const int t_begin = min(warp * per, total), t_end = min(t_begin + per, total);
    int h = t_begin / chunks, ch = t_begin - h * chunks;
    unsigned e_acc = 0;
    int ok_acc = 0;
    const uint8_t* const bgr_al = s.gat.bgr_al;
    const uint32_t* const mask = s.gat.mask;
    const unsigned delta = s.gat.delta;
    // The pose of the current hypothesis is staged in a per-warp shared-memory slot and re-read (broadcast) by
    // every task: nine doubles held in registers would spill.
    double2* const wp = s.wpose[warp];
    auto load_tf = [&](int hh) {
        const Tf* tf = reinterpret_cast<const Tf*>(tf_base + (size_t)(sel ? (int)sel[hh] : hh) * tf_stride);
        __syncwarp();
        if (lane < 3) wp[lane] = make_double2(tf->R[lane * 3], tf->R[lane * 3 + 1]);
        else if (lane == 3) wp[3] = make_double2(tf->t[0], tf->t[1]);
        else if (lane == 4) wp[4] = make_double2(tf->t[2], 0.0);
        else if (!PLANAR && lane == 5) wp[5] = make_double2(tf->R[2], tf->R[5]);
        else if (!PLANAR && lane == 6) wp[6] = make_double2(tf->R[8], 0.0);
        __syncwarp();
    };
    if (t_begin < t_end) load_tf(h);
#ifndef SEARCH_BLK
#define SEARCH_BLK 4
#endif
    constexpr int BLK = SEARCH_BLK;
    constexpr unsigned NONE = 0xffffffffu;
    const int W = a.W, H = a.H;
    for (int t = t_begin; t < t_end;) {
        const int nb = min(BLK, min(chunks - ch, t_end - t));  // warp-uniform
        unsigned key[BLK];  // (y << 16) | x of an in-frame point
#pragma unroll
        for (int b = 0; b < BLK; ++b) {
            key[b] = NONE;
            if (b >= nb) continue;
            const int pi = (ch + b) * 32 + lane;
            double u, v;  // (the map arrays are zero beyond n_map: no clamp of pi)
            bool ok = project_bf<PLANAR, DIST>(wp, s.sf, s.map_x[pi], s.map_y[pi], PLANAR ? 0.0 : s.map_z[pi], s.atan_tab,
                                               a.Wd, a.Hd, u, v);
            ok = ok && pi < n_map;
            const int x = __double2int_rn(u), y = __double2int_rn(v);  // ok implies x >= 0, y >= 0
            const bool inside = ok && x < W && y < H;
            ok_acc += ok ? 1 : 0;
            if (inside) key[b] = ((unsigned)y << 16) | (unsigned)x;
            if (ok && !inside) {  // x == W or y == H: the reference's linear addressing
                rare_point(s, h, x, y);
            }
        }
        // the mask words of the block
        uint32_t mw[BLK];
#pragma unroll
        for (int b = 0; b < BLK; ++b) {
            const unsigned k = key[b];
            mw[b] = 0u;
            if (k != NONE) mw[b] = mask[(k >> 16) * (unsigned)a.WW + ((k & 0xffffu) >> 5)] >> (k & 31u);
        }
        // the pixel words: all ones = no contribution, zero = masked out = black
        uint32_t w0[BLK], w1[BLK], sh8[BLK];
#pragma unroll
        for (int b = 0; b < BLK; ++b) {
            const unsigned k = key[b];
            const unsigned o = (k >> 16) * (unsigned)a.stride + 3u * (k & 0xffffu) + delta;
            w0[b] = w1[b] = k != NONE ? 0u : ~0u;
            sh8[b] = (o & 3u) << 3;
            const bool on = k != NONE && (mw[b] & 1u);
            const uint32_t* p = reinterpret_cast<const uint32_t*>(bgr_al + (o & ~3u));
            if (on) w0[b] = __ldg(p);
            if (on && (o & 2u)) w1[b] = __ldg(p + 1);
        }
#pragma unroll
        for (int b = 0; b < BLK; ++b) {
            const uint32_t inv = ~__funnelshift_r(w0[b], w1[b], sh8[b]) & 0xffffffu;
            e_acc = __dp4a(inv, inv, e_acc);
        }
        ch += nb; t += nb;
        if (ch == chunks || t == t_end) {
            const unsigned e = __reduce_add_sync(0xffffffffu, e_acc);
            const int okw = __reduce_add_sync(0xffffffffu, ok_acc);
            if (lane == 0 && okw) {
                atomicAdd(&s.cnt[h], okw);
                if (e) atomicAdd(&s.sum[h], (unsigned long long)e);
            }
            e_acc = 0; ok_acc = 0;
            if (ch == chunks) { ch = 0; ++h; if (t < t_end) load_tf(h); }
        }
    }
}

// ---- the shift hypotheses (HypothesisGeneration.h:111-140) without their 81 x n_map projections ----------------
// Shift k moves the camera by (sx, sy, 0) in the world: c2w_k p = R^T (p - (t + s_k)) = c2w_best (p - s_k).  On a
// grid map the points p - s_k of all shifts are a small set (2160 distinct points for 81 x 720 on the default map):
// each of them is projected and scored ONCE under the best pose (score_points), and a shift's sums are gathers
// from that table (shift_sums).  The projection of p - s_k differs from the reference's R^T p - R^T (t + s_k) by a few
// ulps of the camera coordinates (~1e-13 px), so a rounding decision can only differ for a coordinate within
// 1e-13 px of a half -- the same class of difference as the fast reciprocal / atan blocks of project_bf, and
// checked the same way (tests/test_parity_full_gpu.py compares all 81 shift errors of every frame with the oracle).
constexpr int SHIFT_PT_CAP = 6399;   // table entries: green_acc + green_xy of `Shared` (not live in this phase)
static_assert(offsetof(Shared, green_xy) == offsetof(Shared, green_acc) + sizeof(unsigned int) * 80 * 40, "the table spans both arrays");
static_assert((SHIFT_PT_CAP + 1) * sizeof(uint32_t) <= 2 * sizeof(unsigned int) * 80 * 40, "table capacity");

// E[q] = 0 for a point that does not project into the frame, else 1 << 31 | its colour error on the masked image
template <bool PLANAR, bool DIST>
__device__ __forceinline__ void score_points(Shared& s, const SearchArgs& a, const Tf* c2w, uint32_t* E)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double2* const wp = s.wpose[warp];
    __syncwarp();
    if (lane < 3) wp[lane] = make_double2(c2w->R[lane * 3], c2w->R[lane * 3 + 1]);
    else if (lane == 3) wp[3] = make_double2(c2w->t[0], c2w->t[1]);
    else if (lane == 4) wp[4] = make_double2(c2w->t[2], 0.0);
    else if (!PLANAR && lane == 5) wp[5] = make_double2(c2w->R[2], c2w->R[5]);
    else if (!PLANAR && lane == 6) wp[6] = make_double2(c2w->R[8], 0.0);
    __syncwarp();
    const int nq = a.n_shift_pts, W = a.W, H = a.H;
    const uint8_t* const bgr_al = s.gat.bgr_al;
    const uint32_t* const mask = s.gat.mask;
    const unsigned delta = s.gat.delta;
    for (int base = warp * 32; base < nq; base += NWARP * 32) {
        const int qi = base + lane;
        double px = 0, py = 0, pz = 0;
        if (qi < nq) { px = __ldg(a.shift_pts + 3 * qi); py = __ldg(a.shift_pts + 3 * qi + 1); pz = __ldg(a.shift_pts + 3 * qi + 2); }
        double u, v;
        bool ok = project_bf<PLANAR, DIST>(wp, s.sf, px, py, PLANAR ? 0.0 : pz, s.atan_tab, a.Wd, a.Hd, u, v);
        ok = ok && qi < nq;
        const int x = __double2int_rn(u), y = __double2int_rn(v);  // ok implies x >= 0, y >= 0
        uint32_t val = 0u;
        if (ok && x < W && y < H) {
            uint32_t w0 = 0u, w1 = 0u;   // zero = masked out = black
            const unsigned o = (unsigned)y * (unsigned)a.stride + 3u * (unsigned)x + delta;
            if ((mask[(unsigned)y * (unsigned)a.WW + ((unsigned)x >> 5)] >> (x & 31)) & 1u) {
                const uint32_t* p = reinterpret_cast<const uint32_t*>(bgr_al + (o & ~3u));
                w0 = __ldg(p);
                if (o & 2u) w1 = __ldg(p + 1);
            }
            const uint32_t inv = ~__funnelshift_r(w0, w1, (o & 3u) << 3) & 0xffffffu;
            val = 0x80000000u | __dp4a(inv, inv, 0u);
        } else if (ok) {  // x == W or y == H: the reference's linear addressing
            val = 0x80000000u | (uint32_t)pixel_err(s.img, x, y, true, 255, 255, 255);
        }
        if (qi < nq) E[qi] = val;
    }
    if (threadIdx.x == 0) E[nq] = 0u;   // what the padding of shift_idx points at
}

// s.sum / s.cnt of the n_shift hypotheses from the table (one warp task = 32 map points of one shift, as in
// score_loop; up to four tasks of one shift at a time so that their index loads are in flight together)
__device__ __forceinline__ void shift_sums(Shared& s, const SearchArgs& a, const uint32_t* E)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int chunks = (a.n_map + 31) >> 5, total = a.n_shift * chunks;
    const int per = (total + NWARP - 1) / NWARP, t_begin = min(warp * per, total), t_end = min(t_begin + per, total);
    int h = t_begin / chunks, ch = t_begin - h * chunks;
    unsigned e_acc = 0, ok_acc = 0;
    for (int t = t_begin; t < t_end;) {
        const int nb = min(4, min(chunks - ch, t_end - t));  // warp-uniform
        const unsigned short* ip = a.shift_idx + (size_t)t * 32 + lane;
        unsigned q[4];
#pragma unroll
        for (int b = 0; b < 4; ++b) q[b] = b < nb ? (unsigned)__ldg(ip + b * 32) : (unsigned)a.n_shift_pts;   // (that entry is empty)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const uint32_t v = E[q[b]];
            e_acc += v & 0x7fffffffu;
            ok_acc += v >> 31;
        }
        ch += nb; t += nb;
        if (ch == chunks || t == t_end) {
            const unsigned e = __reduce_add_sync(0xffffffffu, e_acc), okw = __reduce_add_sync(0xffffffffu, ok_acc);
            if (lane == 0 && okw) {
                atomicAdd(&s.cnt[h], (int)okw);
                if (e) atomicAdd(&s.sum[h], (unsigned long long)e);
            }
            e_acc = 0; ok_acc = 0;
            if (ch == chunks) { ch = 0; ++h; }
        }
    }
}

__global__ void __launch_bounds__(SNT, 1) k_search(SearchArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Shared& s = *reinterpret_cast<Shared*>(smem_raw);
    const int tid = threadIdx.x;
    // map and atan table: once per CTA
    for (int i = tid; i < a.n_map; i += SNT) {
        s.map_x[i] = a.map[3 * i]; s.map_y[i] = a.map[3 * i + 1]; s.map_z[i] = a.map[3 * i + 2];
    }
    for (int i = a.n_map + tid; i < MAX_MAP; i += SNT) { s.map_x[i] = 0; s.map_y[i] = 0; s.map_z[i] = 0; }
    for (int i = tid; i < ATAN_N; i += SNT) s.atan_tab[i] = a.atan_tab[i];
    // persistent CTA: frames are handed out through a ticket counter
    for (;;) {
    __syncthreads();  // the previous frame is done with shared memory
    if (tid == 0) s.frame = (int)atomicAdd(a.ticket, 1u);
    __syncthreads();
    const int frame = s.frame;
    if (frame >= a.n) break;
    const bool ext = a.ext_w2c != nullptr;
    mantis_result* out = ext ? nullptr : a.out + frame;
    mantis_frame_trace* tr = (!ext && a.trace) ? a.trace + frame : nullptr;
    const FrameQuads* fq = ext ? nullptr : a.quads + frame;
    const FramePose* info = ext ? nullptr : a.info + frame;
    const int n_sel = ext ? 0 : info->n_sel;

    if (tid == 0 && !ext) {
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
    if (!ext && (fq->flags || fq->n_quads == 0 || n_sel == 0)) continue;

    if (tid == 0) {
        s.img.bgr = a.bgr + (size_t)frame * a.pitch;
        s.img.mask = a.mask_bits + (size_t)frame * a.H * a.WW;
        s.img.W = a.W; s.img.H = a.H; s.img.WW = a.WW; s.img.stride = a.stride;
    }
    const Img& im = s.img;  // published by the barrier below
    CamK cam;
    {
        const pm::Camera c0 = a.cams[frame];
        cam.fx = c0.fx; cam.fy = c0.fy; cam.cx = c0.cx; cam.cy = c0.cy;
        cam.D0 = c0.D[0]; cam.D1 = c0.D[1]; cam.D2 = c0.D[2]; cam.D3 = c0.D[3];
        cam.has_dist = c0.D[0] != 0 || c0.D[1] != 0 || c0.D[2] != 0 || c0.D[3] != 0;
    }
    const bool planar = a.map_planar != 0;
    if (tid == 32) {
        ScoreFrame& k = s.sf;
        k.fcx = make_double2(cam.fx, cam.cx); k.fcy = make_double2(cam.fy, cam.cy);
        k.D0 = cam.D0; k.D1 = cam.D1; k.D2 = cam.D2; k.D3 = cam.D3;
    }
    const Hyp* hyps = a.hyps + (size_t)frame * MANTIS_MAX_HYPS;
    const short* sel = a.sel + (size_t)frame * MANTIS_MAX_HYPS;
    if (a.mask_in_smem) {
        // The mask plane is gathered once per (hypothesis, point) at 32 unrelated addresses per warp: from global
        // memory that is 32 L1 tag lookups per instruction and the scoring loop ends up bound by the LSU; from
        // shared memory it is a handful of bank conflicts.
        uint32_t* sm = reinterpret_cast<uint32_t*>(smem_raw + sizeof(Shared));
        const uint32_t* gm = a.mask_bits + (size_t)frame * a.H * a.WW;
        for (int i = tid; i < a.H * a.WW; i += SNT) sm[i] = __ldg(gm + i);
    }
    if (tid == 64) {
        const uint8_t* base = a.bgr + (size_t)frame * a.pitch;
        s.gat.delta = (unsigned)((uintptr_t)base & 3u);
        s.gat.bgr_al = base - s.gat.delta;
        s.gat.mask = a.mask_in_smem ? reinterpret_cast<const uint32_t*>(smem_raw + sizeof(Shared))
                                    : a.mask_bits + (size_t)frame * a.H * a.WW;
    }
    long long tk = clock64();
    auto tick = [&](int slot) {
        if (a.dbg && tid == 0) { long long t = clock64(); atomicAdd((unsigned long long*)&a.dbg[slot], (unsigned long long)(t - tk)); tk = t; }
    };

    if (ext) {
        // evaluateHypothesisWithImageWHITE (HypothesisEvaluation.h:107-158) for the caller's poses, through the same
        // scoring loop as the search, 96 poses at a time
        const int mode_e = (planar ? 2 : 0) | (cam.has_dist ? 1 : 0);
        for (int base = 0; base < a.ext_n; base += 96) {
            const int nh = min(96, a.ext_n - base);
            __syncthreads();
            if (tid < nh) {
                const double* src = a.ext_w2c + ((size_t)frame * a.ext_n + base + tid) * 12;
                Tf w;
                for (int k = 0; k < 9; ++k) w.R[k] = src[k];
                for (int k = 0; k < 3; ++k) w.t[k] = src[9 + k];
                s.cand_w2c[tid] = w;
                pm::tf_inv(w, s.cand_c2w[tid]);
            }
            for (int i = tid; i < nh; i += SNT) { s.sum[i] = 0ull; s.cnt[i] = 0; }
            __syncthreads();
            const char* tfb = reinterpret_cast<const char*>(s.cand_c2w);
            if (mode_e == 2) score_loop<true, false>(s, a, nh, tfb, (int)sizeof(Tf), nullptr);
            else if (mode_e == 3) score_loop<true, true>(s, a, nh, tfb, (int)sizeof(Tf), nullptr);
            else if (mode_e == 0) score_loop<false, false>(s, a, nh, tfb, (int)sizeof(Tf), nullptr);
            else score_loop<false, true>(s, a, nh, tfb, (int)sizeof(Tf), nullptr);
            __syncthreads();
            for (int i = tid; i < nh; i += SNT)
                a.ext_scores[(size_t)frame * a.ext_n + base + i] =
                    s.cnt[i] > 0 ? (double)s.sum[i] / ((double)s.cnt[i] * a.projection_bias) : DBLMAX;
        }
        continue;
    }

    // Phases of the search, all through ONE inlined copy of the scoring loop:
    //   0                 (6) evaluateHypotheses on the kept cluster, (7) best 1
    //   1 .. iterations   (8) optimizeHypothesisWithParticleFilter
    //   iterations + 1    (9) the 81 cell shifts, (10) evaluate, (11) best 20
    const int ns = a.n_shift;
    const int keep = ns < 20 ? ns : 20;
    const int mode = (planar ? 2 : 0) | (cam.has_dist ? 1 : 0);  // uniform over the CTA
    for (int ph = 0; ph <= a.iterations + 1; ++ph) {
        int nh;
        const char* tf_base = reinterpret_cast<const char*>(s.cand_c2w);
        int tf_stride = (int)sizeof(Tf);
        const short* selp = nullptr;
        if (ph == 0) {
            nh = n_sel; tf_base = reinterpret_cast<const char*>(&hyps[0].c2w); tf_stride = (int)sizeof(Hyp); selp = sel;
        } else if (ph <= a.iterations) {
            nh = a.particles;
            if (tid < nh) {
                pm::tf_mul(s.best.w2c, a.rand_tf[(ph - 1) * a.particles + tid], s.cand_w2c[tid]);
                pm::tf_inv(s.cand_w2c[tid], s.cand_c2w[tid]);
            }
        } else {
            nh = ns;
            if (tid < ns) {
                Tf w = s.best.w2c;
                w.t[0] += a.shifts[2 * tid];
                w.t[1] += a.shifts[2 * tid + 1];
                w.t[2] += 0;
                s.cand_w2c[tid] = w;
                pm::tf_inv(w, s.cand_c2w[tid]);
            }
        }
        for (int i = tid; i < nh; i += SNT) { s.sum[i] = 0ull; s.cnt[i] = 0; }
        __syncthreads();
        if (ph == a.iterations + 1 && a.n_shift_pts > 0) {
            uint32_t* E = reinterpret_cast<uint32_t*>(s.green_acc);   // green_acc + green_xy: SHIFT_PT_CAP + 1 entries
            if (mode == 2) score_points<true, false>(s, a, &s.best.c2w, E);
            else if (mode == 3) score_points<true, true>(s, a, &s.best.c2w, E);
            else if (mode == 0) score_points<false, false>(s, a, &s.best.c2w, E);
            else score_points<false, true>(s, a, &s.best.c2w, E);
            __syncthreads();
            shift_sums(s, a, E);
        }
        else if (mode == 2) score_loop<true, false>(s, a, nh, tf_base, tf_stride, selp);
        else if (mode == 3) score_loop<true, true>(s, a, nh, tf_base, tf_stride, selp);
        else if (mode == 0) score_loop<false, false>(s, a, nh, tf_base, tf_stride, selp);
        else score_loop<false, true>(s, a, nh, tf_base, tf_stride, selp);
        __syncthreads();
        auto err_of = [&](int i) {
            return s.cnt[i] > 0 ? (double)s.sum[i] / ((double)s.cnt[i] * a.projection_bias) : DBLMAX;
        };
        if (ph >= 1 && ph <= a.iterations) {
            // (8) the greedy update of one refinement iteration, by warp 0: the first particle (lowest index) with the
            // smallest error replaces the best if it is strictly better -- what the sequential `<` scan ends with
            if (tid < 32) {
                double e = DBLMAX;
                int bp = 0x7fffffff;
                for (int p2 = tid; p2 < nh; p2 += 32) {
                    const double ep = err_of(p2);
                    if (ep < e) { e = ep; bp = p2; }
                }
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) {
                    const double eo = __shfl_xor_sync(0xffffffffu, e, d);
                    const int po = __shfl_xor_sync(0xffffffffu, bp, d);
                    if (eo < e || (eo == e && po < bp)) { e = eo; bp = po; }
                }
                if (e < s.best.error) {  // warp-uniform
                    __syncwarp();
                    if (tid < 12) ((double*)&s.best.w2c)[tid] = ((const double*)&s.cand_w2c[bp])[tid];
                    else if (tid < 24) ((double*)&s.best.c2w)[tid - 12] = ((const double*)&s.cand_c2w[bp])[tid - 12];
                    else if (tid == 24) s.best.error = e;
                }
            }
            __syncthreads();
            if (ph == a.iterations) {
                if (tid == 0 && tr) {
                    tr->optimized_error = s.best.error;
                    for (int k = 0; k < 9; ++k) tr->optimized_w2c[k] = s.best.w2c.R[k];
                    for (int k = 0; k < 3; ++k) tr->optimized_w2c[9 + k] = s.best.w2c.t[k];
                }
                tick(1);
            }
            continue;
        }
        for (int i = tid; i < nh; i += SNT) s.err[i] = err_of(i);
        __syncthreads();
        if (ph == 0) {
            if (tid == 0) {
                int bi = 0;  // stable descending sort, last element: minimal error, highest index among ties
                for (int i = 1; i < n_sel; ++i)
                    if (s.err[i] <= s.err[bi]) bi = i;
                s.best = hyps[sel[bi]];
                s.best.error = s.err[bi];
                if (tr) tr->first_best_error = s.err[bi];
            }
            tick(0);
        } else {
            if (tr)
                for (int i = tid; i < ns && i < 81; i += SNT) tr->shift_err[i] = s.err[i];
            // (11) best 20: position in a stable descending sort; keep the last 20.  With 20 shifts or fewer
            // (grid_size <= 4) getBestNHypotheses returns the list as it is (HypothesisEvaluation.h:492), so
            // hyps.back() is the LAST shift, not the best one.
            if (tid < ns) {
                if (ns <= 20) {
                    s.top20[tid] = (short)tid;
                } else {
                    int pos = 0;
                    const double e = s.err[tid];
                    for (int j = 0; j < ns; ++j) pos += (s.err[j] > e) || (s.err[j] == e && j < tid);
                    if (pos >= ns - keep) s.top20[pos - (ns - keep)] = (short)tid;
                }
            }
        }
        __syncthreads();
        if (ph == a.iterations) {  // also when iterations == 0: the refined pose is the first best
            if (tid == 0 && tr) {
                tr->optimized_error = s.best.error;
                for (int k = 0; k < 9; ++k) tr->optimized_w2c[k] = s.best.w2c.R[k];
                for (int k = 0; k < 3; ++k) tr->optimized_w2c[9 + k] = s.best.w2c.t[k];
            }
            tick(1);
        }
    }

    tick(2);
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
    // (a) project every green point of every rotated hypothesis once
    for (int item = tid; item < nrot * ng; item += SNT) {
        const int h = item / ng, gi = item - h * ng, mi = a.n_green_off + gi;
        double u, v;
        const bool ok = project(load_pose(s_rot_c2w[h]), cam, s.map_x[mi], s.map_y[mi], s.map_z[mi], planar, s.atan_tab, im.W, im.H, u, v);
        const int slot = h * 40 + gi;
        s.green_acc[slot] = 0;
        s.green_xy[slot] = ok ? ((__double2int_rn(u) & 0xffff) | (__double2int_rn(v) << 16)) : -1;
        // cvRound(u + dx) == cvRound(u) + dx unless u sits exactly on a half (round-half-even)
        s.green_half[slot] = ok && (u - floor(u) == 0.5 || v - floor(v) == 0.5);
        if (ok && s.green_half[slot]) {
            long long acc = 0;
            for (int dx = -5; dx < 5; ++dx)
                for (int dy = -5; dy < 5; ++dy)
                    acc += pixel_err(im, __double2int_rn(u + (double)dx), __double2int_rn(v + (double)dy), false, 50, 255, 85);
            s.green_acc[slot] = (unsigned int)acc;
        }
    }
    __syncthreads();
    // (b) the 10x10 windows, one window column per work item; integer sums commute
    for (int item = tid; item < nrot * ng * 10; item += SNT) {
        const int slot_i = item / 10, k = item - slot_i * 10;
        const int h = slot_i / ng, gi = slot_i - h * ng, slot = h * 40 + gi;
        const int xy = s.green_xy[slot];
        if (xy < 0 || s.green_half[slot]) continue;
        const int x = (int)(short)(xy & 0xffff) + (k - 5), y0 = (xy >> 16) - 5;
        unsigned int acc = 0;
        if ((unsigned)x < (unsigned)im.W && y0 >= 0 && y0 + 9 < im.H) {
            const uint8_t* p = im.bgr + (size_t)y0 * im.stride + 3 * x;
#pragma unroll
            for (int dy = 0; dy < 10; ++dy, p += im.stride) {
                // the pixel from the two aligned words that hold it; |c - target| per byte, then one IDP.4A
                const uintptr_t q = (uintptr_t)p;
                const uint32_t* w = reinterpret_cast<const uint32_t*>(q & ~(uintptr_t)3);
                const unsigned sh = ((unsigned)q & 3u) << 3;
                const uint32_t w0 = __ldg(w), w1 = sh >= 16u ? __ldg(w + 1) : 0u;
                const uint32_t d = __vabsdiffu4(__funnelshift_r(w0, w1, sh) & 0xffffffu, 0x0055FF32u);  // (B, G, R) = (50, 255, 85)
                acc = __dp4a(d, d, acc);
            }
        } else {
            for (int dy = 0; dy < 10; ++dy) acc += (unsigned)pixel_err(im, x, y0 + dy, false, 50, 255, 85);
        }
        atomicAdd(&s.green_acc[slot], acc);
    }
    tick(3);
    __syncthreads();
    if (tid < nrot) {
        double err = 0;
        int n = 0;
        for (int gi = 0; gi < ng; ++gi)
            if (s.green_xy[tid * 40 + gi] >= 0) { err += (double)s.green_acc[tid * 40 + gi] / 100.0; ++n; }
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
    }  // frame loop
}


// The scoring loop as a function of its own for k_search_rig: compiled apart from the kernel around it (that kernel
// keeps poses and camera models in local memory between calls; inlined into it the loop spilled and mis-scored).
__device__ __noinline__ void score_loop_call(Shared& s, const SearchArgs& a, int nh, int mode)
{
    const char* tfb = reinterpret_cast<const char*>(s.cand_c2w);
    if (mode == 2) score_loop<true, false>(s, a, nh, tfb, (int)sizeof(Tf), nullptr);
    else if (mode == 3) score_loop<true, true>(s, a, nh, tfb, (int)sizeof(Tf), nullptr);
    else if (mode == 0) score_loop<false, false>(s, a, nh, tfb, (int)sizeof(Tf), nullptr);
    else score_loop<false, true>(s, a, nh, tfb, (int)sizeof(Tf), nullptr);
}

// ---- rigs: the same search on the JOINT score of the C cameras of a rig --------------------------------------------
// A hypothesis is the pose of the base; camera k sees it as W2C_k = W2B * b2c_k (include/legacy/mantis2/Mantis2Error.h:
// 67-68), and the colour error and the count of projected points run on over the cameras (the `projected` argument
// of computeImageError, :55-99; MonteCarlo.cpp:233-241 leaves the second camera as a TODO).  One CTA per rig; every
// candidate set goes through score_loop once per camera with the integer sums left to accumulate.  With C == 1 and
// the identity transform this is k_search's arithmetic (tests compare the bytes).
__global__ void __launch_bounds__(SNT, 1) k_search_rig(SearchArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Shared& s = *reinterpret_cast<Shared*>(smem_raw);
    __shared__ Tf s_rot_w2c[80], s_rot_c2w[80];
    __shared__ double s_yaw_err[80];
    __shared__ int s_yaw_n[80];
    __shared__ int s_bi;
    const int tid = threadIdx.x, C = a.C;
    for (int i = tid; i < a.n_map; i += SNT) {
        s.map_x[i] = a.map[3 * i]; s.map_y[i] = a.map[3 * i + 1]; s.map_z[i] = a.map[3 * i + 2];
    }
    for (int i = a.n_map + tid; i < MAX_MAP; i += SNT) { s.map_x[i] = 0; s.map_y[i] = 0; s.map_z[i] = 0; }
    for (int i = tid; i < ATAN_N; i += SNT) s.atan_tab[i] = a.atan_tab[i];
    const bool planar = a.map_planar != 0;
    for (;;) {
    __syncthreads();
    if (tid == 0) s.frame = (int)atomicAdd(a.ticket, 1u);
    __syncthreads();
    const int unit = s.frame;
    if (unit >= a.n) break;
    mantis_result* out = a.out + unit;
    mantis_frame_trace* tr = a.trace ? a.trace + unit : nullptr;
    const FramePose* info = a.info + (size_t)unit * C;
    const int n_sel = info->n_sel;
    int flags = 0, nq = 0, ncont = 0, nraw = 0;
    for (int k = 0; k < C; ++k) {
        const FrameQuads* fq = a.quads + (size_t)unit * C + k;
        flags |= fq->flags; nq += fq->n_quads; ncont += fq->n_contours; nraw += fq->n_quads_raw;
    }
    if (tid == 0) {
        out->n_quads = nq; out->n_hyps = n_sel; out->n_particles = 0; out->error = 0; out->min_yaw_diff = 0;
        for (int k = 0; k < 3; ++k) out->pos[k] = 0;
        for (int k = 0; k < 4; ++k) out->quat_xyzw[k] = 0;
        out->status = flags ? MANTIS_FRAME_OVERFLOW : nq == 0 ? MANTIS_FRAME_NO_QUADS : n_sel == 0 ? MANTIS_FRAME_NO_HYPOTHESES : MANTIS_FRAME_OK;
        if (tr) {
            tr->n_contours = ncont; tr->n_quads_raw = nraw; tr->n_quads = nq; tr->flags = flags;
            tr->n_hyps_generated = info->n_hyps_generated; tr->n_hyps_cluster = n_sel; tr->best_rot = -1; tr->pad_ = 0;
        }
    }
    if (flags || nq == 0 || n_sel == 0) continue;
    const Hyp* hyps = a.hyps + (size_t)unit * C * MANTIS_MAX_HYPS;
    const short* sel = a.sel + (size_t)unit * C * MANTIS_MAX_HYPS;

    // camera k of the rig becomes the one the scoring loop reads; returns its model (registers)
    auto set_camera = [&](int k) -> CamK {
        const int frame = unit * C + k;
        __syncthreads();
        CamK cam;
        const pm::Camera c0 = a.cams[frame];
        cam.fx = c0.fx; cam.fy = c0.fy; cam.cx = c0.cx; cam.cy = c0.cy;
        cam.D0 = c0.D[0]; cam.D1 = c0.D[1]; cam.D2 = c0.D[2]; cam.D3 = c0.D[3];
        cam.has_dist = c0.D[0] != 0 || c0.D[1] != 0 || c0.D[2] != 0 || c0.D[3] != 0;
        if (tid == 0) {
            s.img.bgr = a.bgr + (size_t)frame * a.pitch;
            s.img.mask = a.mask_bits + (size_t)frame * a.H * a.WW;
            s.img.W = a.W; s.img.H = a.H; s.img.WW = a.WW; s.img.stride = a.stride;
        }
        if (tid == 32) {
            ScoreFrame& f = s.sf;
            f.fcx = make_double2(cam.fx, cam.cx); f.fcy = make_double2(cam.fy, cam.cy);
            f.D0 = cam.D0; f.D1 = cam.D1; f.D2 = cam.D2; f.D3 = cam.D3;
        }
        if (tid == 64) {
            const uint8_t* base = a.bgr + (size_t)frame * a.pitch;
            s.gat.delta = (unsigned)((uintptr_t)base & 3u);
            s.gat.bgr_al = base - s.gat.delta;
            s.gat.mask = a.mask_bits + (size_t)frame * a.H * a.WW;   // (global: the planes of C cameras take turns)
        }
        __syncthreads();
        return cam;
    };
    // joint WHITE score of the nh base poses in s.cand_w2c -> s.sum / s.cnt (left for err_of)
    auto score_cands = [&](int nh) {
        for (int i = tid; i < nh; i += SNT) { s.sum[i] = 0ull; s.cnt[i] = 0; }
        for (int k = 0; k < C; ++k) {
            const CamK cam = set_camera(k);
            if (tid < nh) {
                Tf wc;
                pm::tf_mul(s.cand_w2c[tid], a.b2c[unit * C + k], wc);
                pm::tf_inv(wc, s.cand_c2w[tid]);
            }
            __syncthreads();
            score_loop_call(s, a, nh, (planar ? 2 : 0) | (cam.has_dist ? 1 : 0));
        }
        __syncthreads();
    };
    auto err_of = [&](int i) { return s.cnt[i] > 0 ? (double)s.sum[i] / ((double)s.cnt[i] * a.projection_bias) : DBLMAX; };

    // (6) evaluateHypotheses on the kept cluster, 96 at a time, (7) best 1: stable descending sort, last element
    if (tid == 0) { s.best.error = DBLMAX; s_bi = -1; }
    for (int base = 0; base < n_sel; base += 96) {
        const int nh = min(96, n_sel - base);
        __syncthreads();
        if (tid < nh) s.cand_w2c[tid] = hyps[sel[base + tid]].w2c;
        __syncthreads();
        score_cands(nh);
        if (tid == 0)
            for (int i = 0; i < nh; ++i) {
                const double e = err_of(i);
                if (s_bi < 0 || e <= s.best.error) { s.best.error = e; s.best.w2c = s.cand_w2c[i]; s_bi = base + i; }
            }
    }
    __syncthreads();
    if (tid == 0 && tr) tr->first_best_error = s.best.error;

    // (8) optimizeHypothesisWithParticleFilter on the base pose
    for (int it = 0; it < a.iterations; ++it) {
        const int nh = a.particles;
        __syncthreads();
        if (tid < nh) pm::tf_mul(s.best.w2c, a.rand_tf[it * a.particles + tid], s.cand_w2c[tid]);
        __syncthreads();
        score_cands(nh);
        if (tid < 32) {
            double e = DBLMAX;
            int bp = 0x7fffffff;
            for (int p2 = tid; p2 < nh; p2 += 32) {
                const double ep = err_of(p2);
                if (ep < e) { e = ep; bp = p2; }
            }
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) {
                const double eo = __shfl_xor_sync(0xffffffffu, e, d);
                const int po = __shfl_xor_sync(0xffffffffu, bp, d);
                if (eo < e || (eo == e && po < bp)) { e = eo; bp = po; }
            }
            if (e < s.best.error) {
                __syncwarp();
                if (tid < 12) ((double*)&s.best.w2c)[tid] = ((const double*)&s.cand_w2c[bp])[tid];
                else if (tid == 24) s.best.error = e;
            }
        }
        __syncthreads();
    }
    if (tid == 0 && tr) {
        tr->optimized_error = s.best.error;
        for (int k = 0; k < 9; ++k) tr->optimized_w2c[k] = s.best.w2c.R[k];
        for (int k = 0; k < 3; ++k) tr->optimized_w2c[9 + k] = s.best.w2c.t[k];
    }

    // (9) the cell shifts of the base origin, (10) evaluate, (11) best 20
    const int ns = a.n_shift, keep = ns < 20 ? ns : 20;
    __syncthreads();
    if (tid < ns) {
        Tf w = s.best.w2c;
        w.t[0] += a.shifts[2 * tid];
        w.t[1] += a.shifts[2 * tid + 1];
        w.t[2] += 0;
        s.cand_w2c[tid] = w;
    }
    __syncthreads();
    score_cands(ns);
    for (int i = tid; i < ns; i += SNT) s.err[i] = err_of(i);
    __syncthreads();
    if (tr)
        for (int i = tid; i < ns && i < 81; i += SNT) tr->shift_err[i] = s.err[i];
    if (tid < ns) {
        if (ns <= 20) {
            s.top20[tid] = (short)tid;
        } else {
            int pos = 0;
            const double e = s.err[tid];
            for (int j = 0; j < ns; ++j) pos += (s.err[j] > e) || (s.err[j] == e && j < tid);
            if (pos >= ns - keep) s.top20[pos - (ns - keep)] = (short)tid;
        }
    }
    __syncthreads();

    // (12) determineBestYaw on the original images, joint over the cameras
    if (tid < keep) {
        Tf w = s.cand_w2c[s.top20[tid]];
        for (int r = 0; r < 4; ++r) {
            if (r > 0) { Tf t2; pm::tf_mul(a.rotz, w, t2); w = t2; }
            s_rot_w2c[r * keep + tid] = w;
        }
    }
    const int ng = a.n_green, nrot = 4 * keep;
    if (tid < nrot) { s_yaw_err[tid] = 0; s_yaw_n[tid] = 0; }
    for (int k = 0; k < C; ++k) {
        const CamK cam = set_camera(k);
        const Img& im = s.img;
        if (tid < nrot) {
            Tf wc;
            pm::tf_mul(s_rot_w2c[tid], a.b2c[unit * C + k], wc);
            pm::tf_inv(wc, s_rot_c2w[tid]);
        }
        __syncthreads();
        for (int item = tid; item < nrot * ng; item += SNT) {
            const int h = item / ng, gi = item - h * ng, mi = a.n_green_off + gi;
            double u, v;
            const bool ok = project(load_pose(s_rot_c2w[h]), cam, s.map_x[mi], s.map_y[mi], s.map_z[mi], planar, s.atan_tab, im.W, im.H, u, v);
            const int slot = h * 40 + gi;
            unsigned int acc = 0;
            if (ok)
                for (int dx = -5; dx < 5; ++dx)
                    for (int dy = -5; dy < 5; ++dy)
                        acc += (unsigned)pixel_err(im, __double2int_rn(u + (double)dx), __double2int_rn(v + (double)dy), false, 50, 255, 85);
            s.green_acc[slot] = acc;
            s.green_xy[slot] = ok ? 0 : -1;
        }
        __syncthreads();
        if (tid < nrot)
            for (int gi = 0; gi < ng; ++gi)
                if (s.green_xy[tid * 40 + gi] >= 0) { s_yaw_err[tid] += (double)s.green_acc[tid * 40 + gi] / 100.0; ++s_yaw_n[tid]; }
    }
    __syncthreads();
    if (tid < nrot) s.hyp_err[tid] = s_yaw_n[tid] > 0 ? s_yaw_err[tid] / ((double)s_yaw_n[tid] * a.color_bias) : DBLMAX;
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
    }  // rig loop
}

}  // namespace

static int launch_search_impl(mantis_ctx* c, const uint8_t* d_bgr, size_t pitch, int stride, const uint32_t* d_mask_bits, int n,
                              int w, int h, const pm::Camera* d_cams, const FrameQuads* d_quads, const Hyp* d_hyps,
                              const short* d_sel, const FramePose* d_info, mantis_result* d_out, mantis_frame_trace* d_trace,
                              const double* d_ext_w2c, int ext_n, double* d_ext_scores, int rig_cams = 0,
                              const Tf* d_b2c = nullptr)
{
    if (c->n_map <= 0 || c->n_map > MAX_MAP || c->n_green > 40 || c->n_shift > 96 || c->cfg.particles > 96) {
        snprintf(c->err, sizeof(c->err), "map / search sizes exceed the kernel's limits (map %d, green %d)", c->n_map, c->n_green);
        return MANTIS_ERR_INVALID;
    }
    SearchArgs a;
    a.bgr = d_bgr; a.pitch = pitch; a.stride = stride;
    a.mask_bits = d_mask_bits;
    a.W = w; a.H = h; a.WW = words_per_row(w);
    a.Wd = (double)w; a.Hd = (double)h;
    a.cams = d_cams; a.quads = d_quads; a.hyps = d_hyps; a.sel = d_sel; a.info = d_info;
    a.map = (const double*)c->d_map.p;
    a.n_map = c->n_map; a.n_green_off = c->n_map - c->n_green; a.n_green = c->n_green;
    a.map_planar = c->map_planar;
    a.atan_tab = (const double2*)c->d_atan.p;
    a.rand_tf = (const Tf*)c->d_rand.p;
    a.shifts = (const double*)c->d_shifts.p;
    a.n_shift = c->n_shift; a.iterations = c->cfg.iterations; a.particles = c->cfg.particles;
    static const bool direct_shifts = getenv("MANTIS_DIRECT_SHIFTS") != nullptr;   // (for tests: every shift projected on its own)
    a.n_shift_pts = (rig_cams > 0 || direct_shifts) ? 0 : c->n_shift_pts;
    a.shift_pts = (const double*)c->d_shift_pts.p;
    a.shift_idx = (const unsigned short*)c->d_shift_idx.p;
    a.rotz = c->rotz;
    a.projection_bias = c->cfg.projection_bias;
    a.color_bias = c->cfg.color_projection_bias;
    a.out = d_out; a.trace = d_trace;
    a.dbg = (long long*)c->dbg_clocks;
    a.ext_w2c = d_ext_w2c; a.ext_n = ext_n; a.ext_scores = d_ext_scores;
    a.C = rig_cams > 0 ? rig_cams : 1; a.b2c = d_b2c;
    // persistent CTAs (one per SM); frames come from a ticket counter
    const int grid = n < c->sm_count ? n : c->sm_count;
    int rc = ctx_reserve(c, c->search_ticket, sizeof(unsigned));
    if (rc) return rc;
    a.ticket = (unsigned*)c->search_ticket.p;
    a.n = n;
    MCHECK(c, cudaMemsetAsync(a.ticket, 0, sizeof(unsigned), c->stream));
    size_t smem = sizeof(Shared);
    if (rig_cams > 0) {   // n = rigs; the mask planes of the rig's cameras are read from global memory
        a.mask_in_smem = 0;
        MCHECK(c, cudaFuncSetAttribute(k_search_rig, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_search_rig<<<grid, SNT, smem, c->stream>>>(a);
        c->launches++;
        MCHECK(c, cudaGetLastError());
        return MANTIS_OK;
    }
    {
        cudaFuncAttributes fa;
        MCHECK(c, cudaFuncGetAttributes(&fa, k_search));
        const size_t room = 227 * 1024 - fa.sharedSizeBytes - 1024, plane = (size_t)h * a.WW * sizeof(uint32_t);
        a.mask_in_smem = smem + plane <= room;
        if (a.mask_in_smem) smem += plane;
    }
    MCHECK(c, cudaFuncSetAttribute(k_search, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_search<<<grid, SNT, smem, c->stream>>>(a);
    c->launches++;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}

int launch_search(mantis_ctx* c, const uint8_t* d_bgr, size_t pitch, int stride, const uint32_t* d_mask_bits, int n,
                  int w, int h, const pm::Camera* d_cams, const FrameQuads* d_quads, const Hyp* d_hyps,
                  const short* d_sel, const FramePose* d_info, mantis_result* d_out, mantis_frame_trace* d_trace)
{
    return launch_search_impl(c, d_bgr, pitch, stride, d_mask_bits, n, w, h, d_cams, d_quads, d_hyps, d_sel, d_info, d_out,
                              d_trace, nullptr, 0, nullptr);
}

// Score-only launch: WHITE scores of n_hyps caller-supplied poses per frame on the masked image.
int launch_score(mantis_ctx* c, const uint8_t* d_bgr, size_t pitch, int stride, const uint32_t* d_mask_bits, int n, int w,
                 int h, const pm::Camera* d_cams, const double* d_w2c, int n_hyps, double* d_scores)
{
    return launch_search_impl(c, d_bgr, pitch, stride, d_mask_bits, n, w, h, d_cams, nullptr, nullptr, nullptr, nullptr, nullptr,
                              nullptr, d_w2c, n_hyps, d_scores);
}

// Rigs: n_rigs units of rig_cams consecutive frames; d_hyps / d_sel / d_info as launch_pose(..., rig_cams, ...) left them.
int launch_search_rig(mantis_ctx* c, const uint8_t* d_bgr, size_t pitch, int stride, const uint32_t* d_mask_bits, int n_rigs,
                      int rig_cams, int w, int h, const pm::Camera* d_cams, const FrameQuads* d_quads, const Hyp* d_hyps,
                      const short* d_sel, const FramePose* d_info, const Tf* d_b2c, mantis_result* d_out,
                      mantis_frame_trace* d_trace)
{
    return launch_search_impl(c, d_bgr, pitch, stride, d_mask_bits, n_rigs, w, h, d_cams, d_quads, d_hyps, d_sel, d_info, d_out,
                              d_trace, nullptr, 0, nullptr, rig_cams, d_b2c);
}
