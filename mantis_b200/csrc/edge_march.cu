// edge_march.cu -- K1, marching form: fused BGR->gray->blur->Sobel->L1 magnitude->NMS->double threshold
// for frames whose rows are 16-byte aligned and whose width is a multiple of 16 (every BASELINE size);
// other shapes keep the tiled kernel of edge_candidates.cu.  Same integer spec, same outputs (weak /
// strong bit planes), bit-exact against the oracle.
//
//   cv::cvtColor(BGR2GRAY)           include/mantis3/QuadDetection.h:209
//   cv::GaussianBlur(3x3, sigma=3)   :211
//   cv::Canny(T, 3T, 3) up to (not including) hysteresis   :212
//
// Shape.  A warp owns a 256-px column strip (lane = 8 px = 24 bytes of BGR) and walks down a band of rows;
// lanes 1..30 produce output (240 px), lanes 0 and 31 only feed their neighbours.  Only the raw rows
// touch shared memory; every intermediate lives in registers:
//   rows in    one cp.async.bulk (TMA, 1-D) per row into a 3-slot ring per warp, completion on an mbarrier;
//              lanes read their 24 bytes with three conflict-free LDS.64.  A warp keeps up to 3 rows
//              (2.3 KB) in flight, so 16 resident warps cover HBM latency without a register prefetch.
//   gray       two IDP.2A per pixel: the 16-bit operand holds 2 x (3735, 19235, 9798), the byte operand is
//              the raw BGR word, so no byte alignment / extraction is needed; the value lands in byte 2.
//              PRMT packs two pixels into a 16x2 pair.
//   blur       vertical (84, 88, 84) on 16x2 pairs (three-row register window, < 2^16 per half), then
//              horizontal (84, 88, 84) as two IDP.2A per pixel on the pairs (+ neighbours by SHFL);
//              rounding constant and the fp16 exponent 0x64 ride in the accumulator: byte 2 = blurred
//              value b, byte 3 = 0x64, so one PRMT makes half2(1024 + b0, 1024 + b1).
//   Sobel      fp16x2 (HADD2 / HFMA2): every quantity is an integer of magnitude <= 2048, so the
//              arithmetic is exact; the 1024 bias cancels in every difference; |dx| + |dy| is one HADD2
//              with |.| operand modifiers.  Three-row register windows of blurred pairs and their
//              horizontal differences.
//   NMS        branch-free on the same pairs, one row behind the Sobel (three-row register window of
//              magnitudes and their shifted pairs).  OpenCV's fixed-point sector test
//              (|dy| << 15 against 13573 |dx| and (13573 + 65536) |dx|) is evaluated EXACTLY in fp16:
//              13573 = 53 * 256 + 5, so D = 32768 ay - 13573 ax = 256 E - 5 ax with E = 128 ay - 53 ax;
//              F = fma(ax, -53/128, ay) = E / 128 is exact whenever |E| < 2048 and otherwise so large that
//              its sign decides, and G = fma(ax, -5/32768, F) then has the sign of D (a correctly rounded
//              non-zero stays non-zero; subnormals are kept).  Likewise 79109 = 309 * 256 + 5.  The two
//              neighbours to beat are selected with LOP3 masks and compared as half2.
//   out        weak / strong bit planes, one byte (8 px) per lane per row.  No atomics, no lists, no
//              warp synchronisation besides the one that frees a ring slot: cost per row is the same
//              whatever the edge density.
#include <cuda_fp16.h>
#include "ctx.h"

namespace {

constexpr int MW = 4;          // warps per CTA (independent; the CTA is a container)
constexpr int STRIP = 256;     // px per warp, halo lanes included
constexpr int VALID = 240;     // output px per warp
constexpr int NSLOT = 3;       // rows in flight per warp; slot = ring sequence number % 3 = the unroll phase of the march
constexpr int SLOT_BYTES = 800;

struct __align__(16) WarpSmem {
    unsigned char rows[NSLOT][SLOT_BYTES];
    unsigned long long full[NSLOT];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t dp2a_lo(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t dp2a_hi(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("dp2a.hi.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ __half2 h2(uint32_t v) { return *reinterpret_cast<__half2*>(&v); }
__device__ __forceinline__ uint32_t u32(__half2 v) { return *reinterpret_cast<uint32_t*>(&v); }
__device__ __forceinline__ uint32_t hsub(uint32_t a, uint32_t b) { return u32(__hsub2(h2(a), h2(b))); }
__device__ __forceinline__ uint32_t hadd(uint32_t a, uint32_t b) { return u32(__hadd2(h2(a), h2(b))); }
__device__ __forceinline__ uint32_t hmad2(uint32_t a, uint32_t c)  // 2 a + c
{
    return u32(__hfma2(h2(a), __float2half2_rn(2.0f), h2(c)));
}
__device__ __forceinline__ uint32_t habsadd(uint32_t a, uint32_t b) { return u32(__hadd2(__habs2(h2(a)), __habs2(h2(b)))); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!ok);
}

// One lane of a converged warp (elect.sync): the copy engine is driven by a single thread, and with an elected
// lane ptxas emits the R2UR + UBLKCP sequence once instead of a broadcast loop over the active lanes.
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n selp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
    return i;
}

struct MarchArgs {
    const uint8_t* bgr;
    size_t frame_pitch;
    int W, H, stride, WW;
    int n_strips, n_bands, band_h, n_tasks;
    uint32_t low_h2, high_h2;  // thresholds as fp16 bit patterns in both halves
    uint32_t* weak;
    uint32_t* strong;
};

// Registers of a warp's march: three-row windows, slot (i + k) % 3 at step i.
struct Win {
    uint32_t G[3][4];    // gray pairs
    uint32_t B[3][6];    // blurred pairs as half2(1024 + b): [0..3] own, [4] left neighbour's last, [5] right neighbour's first
    uint32_t HD[3][4];   // horizontal differences b(x+1) - b(x-1) as half2
    uint32_t M[3][4];    // |dx| + |dy|
    uint32_t S[3][5];    // shifted magnitude pairs (m(2j-1), m(2j)), j = 0..4
    uint32_t DX[3][4], DY[3][4];
};

struct Task {
    const uint8_t* img;      // frame base
    size_t plane;            // word offset of the frame's bit planes
    int x_base;              // image x of lane 0, pixel 0 (may be -8)
    int y0, y1;              // output rows [y0, y1)
    int a0, nbytes;          // byte range of a row brought in by the bulk copy: [a0, a0 + nbytes)
    int lane_off;            // this lane's 24 bytes inside a slot
    int ybs;                 // first blur row of the march
    bool in_img, left_edge, right_edge, valid_out;
};

// ---- gray pairs of this lane's 8 px from its 24 raw bytes ----
__device__ __forceinline__ void gray_pairs(const unsigned char* slot, int lane_off, uint32_t g[4])
{
    // 2 x (3735, 19235, 9798) as 16-bit operands; 2 x 16384 rounding: the gray value is byte 2, byte 3 is 0
    constexpr uint32_t cB = 7470u, cG = 38470u, cR = 19596u, RND = 32768u;
    constexpr uint32_t A_BG = cB | (cG << 16), A_R0 = cR, A_0B = cB << 16, A_GR = cG | (cR << 16);
    const uint2* sp = reinterpret_cast<const uint2*>(slot + lane_off);
    const uint2 r0 = sp[0], r1 = sp[1], r2 = sp[2];
    const uint32_t w[6] = {r0.x, r0.y, r1.x, r1.y, r2.x, r2.y};
    uint32_t t[8];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const uint32_t w0 = w[3 * q], w1 = w[3 * q + 1], w2 = w[3 * q + 2];
        t[4 * q + 0] = dp2a_hi(A_R0, w0, dp2a_lo(A_BG, w0, RND));
        t[4 * q + 1] = dp2a_lo(A_GR, w1, dp2a_hi(A_0B, w0, RND));
        t[4 * q + 2] = dp2a_lo(A_R0, w2, dp2a_hi(A_BG, w1, RND));
        t[4 * q + 3] = dp2a_hi(A_GR, w2, dp2a_lo(A_0B, w2, RND));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) g[j] = __byte_perm(t[2 * j], t[2 * j + 1], 0x7632);
}

// ---- blur row from the gray window (a, b, c = rows above, at, below), then its horizontal differences ----
__device__ __forceinline__ void blur_row(const uint32_t* Ga, const uint32_t* Gb, const uint32_t* Gc, const Task& T, uint32_t* Bn,
                                         uint32_t* HDn)
{
    uint32_t P[6];  // vertical (84, 88, 84) sums as 16x2 pairs; [4] left neighbour's last pair, [5] right neighbour's first
#pragma unroll
    for (int j = 0; j < 4; ++j) P[j] = (Ga[j] + Gc[j]) * 84u + Gb[j] * 88u;
    P[4] = __shfl_up_sync(0xffffffffu, P[3], 1);
    P[5] = __shfl_down_sync(0xffffffffu, P[0], 1);
    if (T.left_edge) P[4] = P[0];    // BORDER_REFLECT_101: column -1 = column 1 (the high half of pair 0)
    if (T.right_edge) P[5] = P[3];   // column W = column W - 2 (the low half of pair 3)
    // horizontal (84, 88, 84): even px = 84 hi(P[j-1]) + 88 lo(P[j]) + 84 hi(P[j]); odd px = 84 lo(P[j]) + 88 hi(P[j]) + 84 lo(P[j+1])
    constexpr uint32_t KA = 0x54005458u, KB = 0x00545854u, C0 = 0x64008000u;
    uint32_t u[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t pl = j == 0 ? P[4] : P[j - 1], pr = j == 3 ? P[5] : P[j + 1];
        u[2 * j] = dp2a_lo(P[j], KA, dp2a_hi(pl, KA, C0));
        u[2 * j + 1] = dp2a_lo(P[j], KB, dp2a_hi(pr, KB, C0));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) Bn[j] = __byte_perm(u[2 * j], u[2 * j + 1], 0x7632);  // half2(1024 + b, 1024 + b')
    Bn[4] = __shfl_up_sync(0xffffffffu, Bn[3], 1);
    Bn[5] = __shfl_down_sync(0xffffffffu, Bn[0], 1);
    if (T.left_edge) Bn[4] = __byte_perm(Bn[0], 0u, 0x1010);    // BORDER_REPLICATE for the Sobel: b(-1) = b(0)
    if (T.right_edge) Bn[5] = __byte_perm(Bn[3], 0u, 0x3232);   // b(W) = b(W-1)
    uint32_t Q[5];  // (b(2j-1), b(2j))
    Q[0] = __byte_perm(Bn[4], Bn[0], 0x5432);
#pragma unroll
    for (int j = 1; j < 4; ++j) Q[j] = __byte_perm(Bn[j - 1], Bn[j], 0x5432);
    Q[4] = __byte_perm(Bn[3], Bn[5], 0x5432);
#pragma unroll
    for (int j = 0; j < 4; ++j) HDn[j] = hsub(Q[j + 1], Q[j]);
}

// ---- Sobel at the window's centre row: dx, dy, |dx| + |dy| as half2 pairs ----
__device__ __forceinline__ void sobel_row(const uint32_t* Bt, const uint32_t* Bb, const uint32_t* HDt, const uint32_t* HDc,
                                          const uint32_t* HDb, uint32_t* dx, uint32_t* dy, uint32_t* mg)
{
    uint32_t VD[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) VD[j] = hsub(Bb[j], Bt[j]);   // b(y+1) - b(y-1)
    uint32_t QV[5];
    QV[0] = __byte_perm(VD[4], VD[0], 0x5432);
#pragma unroll
    for (int j = 1; j < 4; ++j) QV[j] = __byte_perm(VD[j - 1], VD[j], 0x5432);
    QV[4] = __byte_perm(VD[3], VD[5], 0x5432);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        dy[j] = hadd(hmad2(VD[j], QV[j]), QV[j + 1]);
        dx[j] = hadd(hmad2(HDc[j], HDt[j]), HDb[j]);
        mg[j] = habsadd(dx[j], dy[j]);
    }
}

// ---- shifted pairs of a magnitude row (neighbours' edge pixels by shuffle) ----
__device__ __forceinline__ void shifted_pairs(const uint32_t* M, uint32_t* S)
{
    const uint32_t ml = __shfl_up_sync(0xffffffffu, M[3], 1), mr = __shfl_down_sync(0xffffffffu, M[0], 1);
    S[0] = __byte_perm(ml, M[0], 0x5432);
#pragma unroll
    for (int j = 1; j < 4; ++j) S[j] = __byte_perm(M[j - 1], M[j], 0x5432);
    S[4] = __byte_perm(M[3], mr, 0x5432);
}

__device__ __forceinline__ uint32_t sel(uint32_t mask, uint32_t a, uint32_t b) { return (a & mask) | (b & ~mask); }
// 0xFFFF in each half whose sign bit is set (prmt's msb-replication mode; __byte_perm ignores that selector bit)
__device__ __forceinline__ uint32_t sign_mask2(uint32_t x)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %1, 0xBB99;" : "=r"(d) : "r"(x));
    return d;
}

// ---- non-max suppression + double threshold of one row, branch-free on half2 pairs ----
// (rows u / c / d = above / at / below; returns the weak and strong bits of this lane's 8 px)
__device__ __forceinline__ void nms_row(const uint32_t* Mu, const uint32_t* Su, const uint32_t* Mc, const uint32_t* Sc,
                                        const uint32_t* Md, const uint32_t* Sd, const uint32_t* DX, const uint32_t* DY,
                                        uint32_t low_h2, uint32_t high_h2, uint32_t& weak8, uint32_t& strong8)
{
    const __half2 k53 = __floats2half2_rn(-0.4140625f, -0.4140625f);        // -53/128
    const __half2 k309 = __floats2half2_rn(-2.4140625f, -2.4140625f);       // -309/128
    const __half2 k5 = __floats2half2_rn(-0.000152587890625f, -0.000152587890625f);  // -5/32768
    const __half2 zero = __floats2half2_rn(0.f, 0.f);
    uint32_t wacc = 0u, sacc = 0u;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const __half2 ax = __habs2(h2(DX[j])), ay = __habs2(h2(DY[j]));
        const __half2 g1 = __hfma2(ax, k5, __hfma2(ax, k53, ay));     // sign of (ay << 15) - 13573 ax
        const __half2 g2 = __hfma2(ax, k5, __hfma2(ax, k309, ay));    // sign of (ay << 15) - 79109 ax
        const uint32_t horiz = __hlt2_mask(g1, zero), vert = __hgt2_mask(g2, zero);
        const uint32_t x = DX[j] ^ DY[j];
        const uint32_t neg = sign_mask2(x);                            // signs differ: 0xFFFF per half
        const uint32_t n1 = sel(horiz, Sc[j], sel(vert, Mu[j], sel(neg, Su[j + 1], Su[j])));
        const uint32_t n2 = sel(horiz, Sc[j + 1], sel(vert, Md[j], sel(neg, Sd[j], Sd[j + 1])));
        // horizontal / vertical: m > n1 && m >= n2; diagonal: both strict (m >= n2 + 1, all integers)
        const uint32_t one = ~(horiz | vert) & 0x3C003C00u;
        const __half2 m = h2(Mc[j]);
        const uint32_t c = __hgt2_mask(m, h2(n1)) & __hge2_mask(m, __hadd2(h2(n2), h2(one))) & __hgt2_mask(m, h2(low_h2));
        const uint32_t st = c & __hgt2_mask(m, h2(high_h2));
        const uint32_t K = (1u << (2 * j)) | (1u << (2 * j + 17));
        wacc |= c & K;
        sacc |= st & K;
    }
    weak8 = (wacc | (wacc >> 16)) & 0xffu;
    strong8 = (sacc | (sacc >> 16)) & 0xffu;
}

// Bulk copy number q of a warp's march: gray row ybs - 1 + q (reflected into the image) into slot q % NSLOT.
// Rows nobody will wait for are never requested (a CTA must not exit with copies in flight).  Prologue only;
// the steps advance a running row pointer instead (march_step).
__device__ __forceinline__ void ring_issue(const MarchArgs& a, const Task& T, WarpSmem& s, int q, int n_steps)
{
    const int step = q - 2;
    if (q >= 2 && (step >= n_steps || T.ybs + step >= a.H)) return;
    const int slot = q % NSLOT;
    mbar_expect_tx(&s.full[slot], (uint32_t)T.nbytes);
    bulk_g2s(s.rows[slot], T.img + (size_t)reflect101(T.ybs - 1 + q, a.H) * a.stride + T.a0, (uint32_t)T.nbytes, &s.full[slot]);
}

// One step of the march: blur row yb (if it exists) enters the windows, the Sobel runs at yc = yb - 1, the NMS at
// yn = yb - 2.  PH = step index mod 3 names the window slots, so that nothing is moved between registers.
template <int PH, bool BORDER>
__device__ __forceinline__ void march_step(const MarchArgs& a, const Task& T, WarpSmem& s, Win& w, int step, int n_steps, int lane,
                                           uint32_t ph, const uint8_t* next_row)
{
    constexpr int SA = PH % 3, SB = (PH + 1) % 3, SC = (PH + 2) % 3;   // oldest, middle, newest
    const int H = a.H, yb = T.ybs + step;
    if (yb < H) {
        // ring sequence number q = step + 2 (q = 0, 1 are the prologue's rows): slot q % 3 = SC, phase parity (q / 3) & 1
        mbar_wait(&s.full[SC], ph ^ (PH > 0 ? 1u : 0u));
        gray_pairs(s.rows[SC], T.lane_off, w.G[SC]);
        __syncwarp();   // every lane has its bytes: the slot may be refilled
        // ... with gray row yb + 4, the one step + 3 consumes -- if that step exists and has a blur row (yb + 3 < H);
        // yb + 4 == H is the reflected row H - 2
        if (step + 3 < n_steps && yb + 3 < H) {
            if (elect_one()) {
                const uint8_t* src = next_row - (yb + 4 == H ? 2 * (size_t)a.stride : 0);
                mbar_expect_tx(&s.full[SC], (uint32_t)T.nbytes);
                bulk_g2s(s.rows[SC], src, (uint32_t)T.nbytes, &s.full[SC]);
            }
        }
        blur_row(w.G[SA], w.G[SB], w.G[SC], T, w.B[SC], w.HD[SC]);
    }
    // rows of this step: yc = yb - 1 gets its Sobel (slot SB), yn = yb - 2 its NMS (slot SA; row yn - 1 sits in SC)
    const int yc = yb - 1;
    if (yc < T.y0 - 1) return;   // the windows are still filling (bands that do not start at the top)
    if (yc >= 0 && yc < H) {
        if (BORDER) {
            // BORDER_REPLICATE of the blurred image above row 0 / below row H-1: the dead window slot takes a copy
            if (yc == 0) {
#pragma unroll
                for (int j = 0; j < 6; ++j) w.B[SA][j] = w.B[SB][j];
#pragma unroll
                for (int j = 0; j < 4; ++j) w.HD[SA][j] = w.HD[SB][j];
            }
            if (yc == H - 1) {
#pragma unroll
                for (int j = 0; j < 6; ++j) w.B[SC][j] = w.B[SB][j];
#pragma unroll
                for (int j = 0; j < 4; ++j) w.HD[SC][j] = w.HD[SB][j];
            }
        }
        sobel_row(w.B[SA], w.B[SC], w.HD[SA], w.HD[SB], w.HD[SC], w.DX[SB], w.DY[SB], w.M[SB]);
        if (!T.in_img) {   // magnitude outside the image is 0
#pragma unroll
            for (int j = 0; j < 4; ++j) w.M[SB][j] = 0u;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) w.M[SB][j] = 0u;
    }
    shifted_pairs(w.M[SB], w.S[SB]);
    const int yn = yc - 1;
    if (yn >= T.y0 && yn < T.y1) {
        uint32_t wk, st;
        nms_row(w.M[SC], w.S[SC], w.M[SA], w.S[SA], w.M[SB], w.S[SB], w.DX[SA], w.DY[SA], a.low_h2, a.high_h2, wk, st);
        if (T.valid_out) {
            const size_t off = (T.plane + (size_t)yn * a.WW) * 4 + (size_t)((T.x_base >> 3) + lane);
            reinterpret_cast<uint8_t*>(a.weak)[off] = (uint8_t)wk;
            reinterpret_cast<uint8_t*>(a.strong)[off] = (uint8_t)st;
        }
    }
}

__global__ void __launch_bounds__(MW * 32, 4) k_edge_march(MarchArgs a)
{
    __shared__ WarpSmem sm[MW];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpSmem& s = sm[warp];
    if (lane == 0) {
        for (int k = 0; k < NSLOT; ++k) mbar_init(&s.full[k], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    const int task = blockIdx.x * MW + warp;
    if (task >= a.n_tasks) return;
    const int strip = task % a.n_strips, band = (task / a.n_strips) % a.n_bands, frame = task / (a.n_strips * a.n_bands);

    Task T;
    T.img = a.bgr + (size_t)frame * a.frame_pitch;
    T.plane = (size_t)frame * a.H * a.WW;
    T.x_base = strip * VALID - 8;
    T.y0 = band * a.band_h;
    T.y1 = min(T.y0 + a.band_h, a.H);
    const int x0 = T.x_base + 8 * lane;
    T.in_img = x0 >= 0 && x0 + 8 <= a.W;
    T.left_edge = x0 == 0;
    T.right_edge = x0 + 8 == a.W;
    T.valid_out = T.in_img && lane >= 1 && lane <= 30;
    const int px_lo = max(T.x_base, 0), px_hi = min(T.x_base + STRIP, a.W);
    T.a0 = (3 * px_lo) & ~15;
    T.nbytes = ((3 * px_hi + 15) & ~15) - T.a0;   // 3 W is a multiple of 16, so this never passes the end of the row
    T.lane_off = 3 * min(max(x0, 0), a.W - 8) - T.a0;

    // blur rows yb = ybs .. y1 + 1 (those inside the image are computed), one per step; Sobel centres lag by one row
    T.ybs = max(T.y0 - 2, 0);
    const int n_steps = T.y1 - T.ybs + 2;
    Win w;
    if (lane == 0)
        for (int q = 0; q < NSLOT; ++q) ring_issue(a, T, s, q, n_steps);
    mbar_wait(&s.full[0], 0u);
    gray_pairs(s.rows[0], T.lane_off, w.G[0]);   // gray row ybs - 1
    mbar_wait(&s.full[1], 0u);
    gray_pairs(s.rows[1], T.lane_off, w.G[1]);   // gray row ybs
    __syncwarp();
    if (lane == 0)
        for (int q = NSLOT; q < NSLOT + 2; ++q) ring_issue(a, T, s, q, n_steps);
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 6; ++j) w.B[0][j] = w.B[1][j] = w.B[2][j] = 0u;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        w.HD[0][j] = w.HD[1][j] = w.HD[2][j] = 0u;
        w.M[0][j] = w.M[1][j] = w.M[2][j] = 0u;
        w.DX[0][j] = w.DX[1][j] = w.DX[2][j] = 0u;
        w.DY[0][j] = w.DY[1][j] = w.DY[2][j] = 0u;
    }
#pragma unroll
    for (int j = 0; j < 5; ++j) w.S[0][j] = w.S[1][j] = w.S[2][j] = 0u;
    // Steps whose Sobel centre is row 0 or row H - 1 take the variant with the border copies; every other group
    // of three runs without them.  (Sobel centre of step k is ybs + k - 1.)
    // next_row: address of gray row yb + 4 for the step at hand (the row its refill asks for)
    const uint8_t* next_row = T.img + (size_t)(T.ybs + 4) * a.stride + T.a0;
    const size_t s1 = (size_t)a.stride;
    uint32_t ph = 0u;
    for (int step = 0; step < n_steps; step += 3, ph ^= 1u, next_row += 3 * s1) {
        const int yc0 = T.ybs + step - 1;
        const bool edge = yc0 <= 0 || yc0 + 2 >= a.H - 1;
        if (edge) {
            march_step<0, true>(a, T, s, w, step, n_steps, lane, ph, next_row);
            if (step + 1 < n_steps) march_step<1, true>(a, T, s, w, step + 1, n_steps, lane, ph, next_row + s1);
            if (step + 2 < n_steps) march_step<2, true>(a, T, s, w, step + 2, n_steps, lane, ph, next_row + 2 * s1);
        } else {
            march_step<0, false>(a, T, s, w, step, n_steps, lane, ph, next_row);
            if (step + 1 < n_steps) march_step<1, false>(a, T, s, w, step + 1, n_steps, lane, ph, next_row + s1);
            if (step + 2 < n_steps) march_step<2, false>(a, T, s, w, step + 2, n_steps, lane, ph, next_row + 2 * s1);
        }
    }
}

}  // namespace

bool edge_march_supported(const uint8_t* d_bgr, int w, int h, int stride, size_t frame_pitch)
{
    return w % 16 == 0 && w >= 16 && h >= 4 && (((uintptr_t)d_bgr | (uintptr_t)stride | (uintptr_t)frame_pitch) & 15u) == 0;
}

int launch_edge_march(mantis_ctx* c, const uint8_t* d_bgr, int n, int w, int h, int stride, size_t frame_pitch, int low, int high,
                      uint32_t* d_weak, uint32_t* d_strong)
{
    MarchArgs a;
    a.bgr = d_bgr; a.frame_pitch = frame_pitch; a.W = w; a.H = h; a.stride = stride; a.WW = words_per_row(w);
    a.n_strips = (w + VALID - 1) / VALID;
    // enough warps to fill the machine a few times over, bands of at least 32 rows
    const long long want = (long long)c->sm_count * 16 * 2;
    long long nb = (want + (long long)n * a.n_strips - 1) / ((long long)n * a.n_strips);
    const int max_bands = (h + 31) / 32;
    if (nb < 1) nb = 1;
    if (nb > max_bands) nb = max_bands;
    a.band_h = (h + (int)nb - 1) / (int)nb;
    a.n_bands = (h + a.band_h - 1) / a.band_h;
    const long long tasks = (long long)n * a.n_strips * a.n_bands;
    if (tasks > 0x7fffffffLL) return MANTIS_ERR_INVALID;
    a.n_tasks = (int)tasks;
    const int lo = low < 2047 ? (low < 0 ? 0 : low) : 2047, hi = high < 2047 ? (high < 0 ? 0 : high) : 2047;
    const __half hl = __float2half_rn((float)lo), hh = __float2half_rn((float)hi);
    const uint16_t lb = *reinterpret_cast<const uint16_t*>(&hl), hb = *reinterpret_cast<const uint16_t*>(&hh);
    a.low_h2 = (uint32_t)lb | ((uint32_t)lb << 16);
    a.high_h2 = (uint32_t)hb | ((uint32_t)hb << 16);
    a.weak = d_weak; a.strong = d_strong;
    k_edge_march<<<(a.n_tasks + MW - 1) / MW, MW * 32, 0, c->stream>>>(a);
    c->launches++;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}
