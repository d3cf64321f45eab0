// edge_link.cu -- K2: per-frame on-chip Canny hysteresis closure, dilate x2 / erode x1, and the
// cleanImageByEdge mask chain, all on bit-packed planes that never leave shared memory.
//
//   cv::Canny's 8-connected hysteresis closure                 QuadDetection.h:212
//   cv::dilate(iter 2) / cv::erode(iter 1)                     QuadDetection.h:213-214
//   cleanImageByEdge: MORPH_GRADIENT (3x3 cross), NOT, findContours(RETR_LIST)+drawContours
//   (== boundary pixels of the NOT image), OR with Canny, closes of 3/4/5 iterations,
//   erode 3                                                    HypothesisEvaluation.h:336-372
//
// One thread-block cluster per frame: the rows of the frame are split across the CTAs of the
// cluster and halo rows are exchanged through distributed shared memory (DSMEM).
//   closure : E <- strong;  repeat E <- weak & flood(E | dilate3x3(E)) until stable.  In-word
//             horizontal runs are filled in O(1) with the carry trick ((M+S)^M)&M.
//   k x (3x3 max) == (2k+1)^2 box OR with 0 outside the image, done as a row pass and a column pass
//   k x (3x3 min) == the same on the complement (1 outside the image)
#include <cooperative_groups.h>
#include "ctx.h"

namespace cg = cooperative_groups;

namespace {

constexpr int LNT = 512;
constexpr int PADR = 8;  // zero rows kept above local row 1 and below local row HC in every plane (largest k of the chain)

// fill every run of M that contains a bit of S (S subset of M)
__device__ __forceinline__ uint32_t flood_word(uint32_t M, uint32_t S)
{
    uint32_t up = ((M + S) ^ M) & M;
    uint32_t Mr = __brev(M), Sr = __brev(S);
    uint32_t dn = __brev(((Mr + Sr) ^ Mr) & Mr);
    return up | dn | S;
}

__device__ __forceinline__ uint32_t expand4(uint32_t nib)
{
    return ((nib * 0x00204081u) & 0x01010101u) * 0xFFu;
}

struct LinkArgs {
    const uint32_t* weak;
    const uint32_t* strong;
    uint32_t* canny_bits;  // required when mask output is requested
    uint32_t* morph_bits;
    uint32_t* mask_bits;
    uint8_t* canny_u8;
    uint8_t* morph_u8;
    uint8_t* mask_u8;
    int W, H, WW, HC;  // HC = rows per CTA
    long long* dbg;    // tuning aid: clock64 per phase (null = off)
};

// Geometry of this CTA's band of rows.  Planes are stored as (HC+2) x WW words: local row 0
// and nrows+1 are halo rows (copies of the neighbouring CTAs' boundary rows).
struct Band {
    int W, H, WW, HC, rank, last_rank, row0, nrows;
    uint32_t last_mask;
};

// write rows of a bit plane held in smem (local rows 1..nrows) as u8 {0,255}
__device__ void store_u8(const uint32_t* s, const Band& b, uint8_t* dst)
{
    const int W = b.W, WW = b.WW;
    if ((W & 15) == 0 && (((uintptr_t)dst) & 15) == 0) {
        int halves = W >> 4;  // 16-px groups per row
        for (int i = threadIdx.x; i < b.nrows * halves; i += LNT) {
            int r = i / halves, hx = i - r * halves;
            uint32_t w = s[(r + 1) * WW + (hx >> 1)];
            uint32_t v16 = (hx & 1) ? (w >> 16) : (w & 0xffffu);
            uint4 v;
            v.x = expand4(v16 & 15u);
            v.y = expand4((v16 >> 4) & 15u);
            v.z = expand4((v16 >> 8) & 15u);
            v.w = expand4((v16 >> 12) & 15u);
            *reinterpret_cast<uint4*>(dst + (size_t)(b.row0 + r) * W + hx * 16) = v;
        }
    } else {
        for (int i = threadIdx.x; i < b.nrows * W; i += LNT) {
            int r = i / W, x = i - r * W;
            uint32_t w = s[(r + 1) * WW + (x >> 5)];
            dst[(size_t)(b.row0 + r) * W + x] = ((w >> (x & 31)) & 1u) ? 255 : 0;
        }
    }
}

__device__ void store_bits(const uint32_t* s, const Band& b, uint32_t* dst_plane)
{
    for (int i = threadIdx.x; i < b.nrows * b.WW; i += LNT) dst_plane[(size_t)b.row0 * b.WW + i] = s[b.WW + i];
}

// Pull the neighbours' boundary rows of plane `s` (same smem offset in every CTA) into my halo
// rows.  Returns (per thread) whether a halo word changed.  Rows outside the image are never
// read through the halo (see row_word), so nothing is written for them.
__device__ int pull_halo(cg::cluster_group& cluster, uint32_t* s, const Band& b)
{
    int changed = 0;
    const int WW = b.WW;
    for (int j = threadIdx.x; j < 2 * WW; j += LNT) {
        int top = j < WW;
        int jj = top ? j : j - WW;
        uint32_t v;
        uint32_t* dst;
        if (top) {
            if (b.rank == 0) continue;
            const uint32_t* nb = cluster.map_shared_rank(s, b.rank - 1);
            v = nb[b.HC * WW + jj];  // every non-last CTA owns exactly HC rows
            dst = &s[jj];
        } else {
            if (b.rank >= b.last_rank) continue;
            const uint32_t* nb = cluster.map_shared_rank(s, b.rank + 1);
            v = nb[1 * WW + jj];
            dst = &s[(b.nrows + 1) * WW + jj];
        }
        if (*dst != v) { *dst = v; changed = 1; }
    }
    return changed;
}

// Make the halo rows of `s` current: cluster barrier (everyone finished writing), pull, block
// barrier.  For a single-CTA frame only the block barrier is needed.
__device__ void refresh_halo(cg::cluster_group& cluster, uint32_t* s, const Band& b, int CS)
{
    if (CS > 1) {
        cluster.sync();
        pull_halo(cluster, s, b);
    }
    __syncthreads();
}

// word (local row lr in 0..nrows+1, column j) of a plane, with `fill` for anything outside the
// image (rows above 0 / below H-1, columns left of 0 / right of W-1 including the unused
// bits of the last word).
__device__ __forceinline__ uint32_t row_word(const uint32_t* s, const Band& b, int lr, int j, uint32_t fill)
{
    int gr = b.row0 + lr - 1;
    if (gr < 0 || gr >= b.H || j < 0 || j >= b.WW) return fill;
    uint32_t w = s[lr * b.WW + j];
    if (j == b.WW - 1) w |= (fill & ~b.last_mask);
    return w;
}

// 3-row vertical combine + horizontal neighbours for one output word.
// FULL: 3x3 box; otherwise the 3x3 cross.  OR: max with 0 outside; otherwise min with 1 outside.
template <bool OR, bool FULL>
__device__ __forceinline__ uint32_t stencil(const uint32_t* s, const Band& b, int lr, int j)
{
    const uint32_t fill = OR ? 0u : 0xffffffffu;
    uint32_t u = row_word(s, b, lr - 1, j, fill), c = row_word(s, b, lr, j, fill), d = row_word(s, b, lr + 1, j, fill);
    uint32_t cl = row_word(s, b, lr, j - 1, fill), cr = row_word(s, b, lr, j + 1, fill);
    uint32_t v, vl, vr;  // columns whose left / right neighbours are taken
    if (FULL) {
        uint32_t ul = row_word(s, b, lr - 1, j - 1, fill), ur = row_word(s, b, lr - 1, j + 1, fill);
        uint32_t dl = row_word(s, b, lr + 1, j - 1, fill), dr = row_word(s, b, lr + 1, j + 1, fill);
        v = OR ? (u | c | d) : (u & c & d);
        vl = OR ? (ul | cl | dl) : (ul & cl & dl);
        vr = OR ? (ur | cr | dr) : (ur & cr & dr);
    } else {
        v = c; vl = cl; vr = cr;
    }
    uint32_t left = (v << 1) | (vl >> 31), right = (v >> 1) | (vr << 31);
    uint32_t o;
    if (FULL) o = OR ? (v | left | right) : (v & left & right);
    else o = OR ? (u | c | d | left | right) : (u & c & d & left & right);
    if (j == b.WW - 1) o &= b.last_mask;
    return o;
}

// k iterations of the 3x3 dilation (erode = false) or erosion (erode = true) in two separable passes:
// src -(rows: OR of shifts -k..k)-> tmp -(columns: OR of rows -k..k)-> src.  Erosion runs on the
// complement (outside the image counts as 1, i.e. never wins the minimum).  One compact routine for every
// morphological step of the kernel keeps the code inside the instruction cache.
__device__ __noinline__ void morph_op(cg::cluster_group& cluster, uint32_t* src, uint32_t* tmp, const Band& b, int k,
                                      bool erode, int CS)
{
    const int WW = b.WW, n = b.nrows * WW;
    const uint32_t inv = erode ? 0xffffffffu : 0u;
    // (row, word) of this thread's first item and the step of LNT items, so that the loops need no division
    const int lr0 = threadIdx.x / WW + 1, j0 = threadIdx.x - (lr0 - 1) * WW, dlr = LNT / WW, dj = LNT - dlr * WW;
    int lr = lr0, j = j0;
    for (int i = threadIdx.x; i < n; i += LNT) {
        const uint32_t* row = src + lr * WW;
        uint32_t C = row[j] ^ inv, L = j > 0 ? row[j - 1] ^ inv : 0u, R = j + 1 < WW ? row[j + 1] ^ inv : 0u;
        if (j == WW - 1) C &= b.last_mask;
        if (j + 1 == WW - 1) R &= b.last_mask;
        uint32_t acc = C;
        for (int d = 1; d <= k; ++d) acc |= __funnelshift_l(L, C, d) | __funnelshift_r(C, R, d);
        if (j == WW - 1) acc &= b.last_mask;
        tmp[lr * WW + j] = acc;
        lr += dlr; j += dj;
        if (j >= WW) { j -= WW; ++lr; }
    }
    if (CS > 1) cluster.sync(); else __syncthreads();
    // column pass: rows of this CTA's band from local shared memory, rows of the neighbouring bands (only
    // when the frame is split over a cluster) straight from their shared memory; rows outside the image add 0
    const uint32_t* up = nullptr;
    const uint32_t* down = nullptr;
    int down_rows = 0;
    if (CS > 1) {
        if (b.rank > 0 && b.nrows > 0) up = cluster.map_shared_rank(tmp, b.rank - 1);
        if (b.rank < b.last_rank) {
            down = cluster.map_shared_rank(tmp, b.rank + 1);
            down_rows = min(b.HC, b.H - (b.rank + 1) * b.HC);
        }
    }
    lr = lr0; j = j0;
    for (int i = threadIdx.x; i < n; i += LNT) {
        const int lo = max(1, lr - k), hi = min(b.nrows, lr + k);
        uint32_t acc = 0;
        for (int r = lo; r <= hi; ++r) acc |= tmp[r * WW + j];
        if (up)
            for (int r = lr - k; r < 1; ++r) acc |= up[(b.HC + r) * WW + j];
        if (down)
            for (int r = b.nrows + 1; r <= min(lr + k, b.nrows + down_rows); ++r) acc |= down[(r - b.nrows) * WW + j];
        acc ^= inv;
        if (j == WW - 1) acc &= b.last_mask;
        src[lr * WW + j] = acc;
        lr += dlr; j += dj;
        if (j >= WW) { j -= WW; ++lr; }
    }
    if (CS > 1) cluster.sync(); else __syncthreads();
}

// Same operation for a frame held by ONE CTA (the common case up to 640x480), with k and the row length known
// at compile time.  A thread keeps one word column j and visits rows lr = 1 + tid / WW, + LNT / WW, ... so
// that everything that depends on j (neighbour masks, the last word's mask) is loop-invariant, and the
// column pass reads its 2k+1 rows at immediate offsets; rows outside the image are the zero rows every plane
// carries (PADR of them), so neither pass tests a bound.  3k+7 instructions per word in the row pass, about k+6 in
// the column pass (see there).
template <int K, int WWC>
__device__ __forceinline__ void morph_fast_t(cg::cluster_group& cluster, uint32_t* src, uint32_t* tmp, const Band& b, bool erode,
                                             int CS)
{
    const int WW = WWC ? WWC : b.WW;
    const int R = LNT / WW;
    const int tid = threadIdx.x;
    const bool active = tid < R * WW;
    const int rb = tid / WW, j = tid - rb * WW;
    const uint32_t inv = erode ? 0xffffffffu : 0u;
    const uint32_t mC = (j == WW - 1) ? b.last_mask : 0xffffffffu;
    const uint32_t mL = j > 0 ? 0xffffffffu : 0u;
    const uint32_t mR = (j + 1 < WW) ? ((j + 1 == WW - 1) ? b.last_mask : 0xffffffffu) : 0u;
    if (active) {
        for (int lr = 1 + rb; lr <= b.nrows; lr += R) {
            const uint32_t* row = src + lr * WW + j;  // row[-1] / row[1] of the first / last word exist (halo rows) and are masked
            const uint32_t C = (row[0] ^ inv) & mC, L = (row[-1] ^ inv) & mL, Rw = (row[1] ^ inv) & mR;
            uint32_t acc = C;
#pragma unroll
            for (int d = 1; d <= K; ++d) acc |= __funnelshift_l(L, C, d) | __funnelshift_r(C, Rw, d);
            tmp[lr * WW + j] = acc & mC;
        }
    }
    if (CS > 1) {
        // a frame split over a cluster: the K rows on either side of the band come from the neighbours' planes
        // (DSMEM) into this plane's pad rows, so that the column pass below stays local and branch-free
        cluster.sync();  // every CTA has written its rows of tmp
        const uint32_t* up = (b.rank > 0 && b.nrows > 0) ? cluster.map_shared_rank(tmp, b.rank - 1) : nullptr;
        const uint32_t* dn = b.rank < b.last_rank ? cluster.map_shared_rank(tmp, b.rank + 1) : nullptr;
        const int down_rows = dn ? min(b.HC, b.H - (b.rank + 1) * b.HC) : 0;
        for (int i = tid; i < K * WW; i += LNT) {
            const int r = i / WW, jj = i - r * WW;
            if (up) tmp[-r * WW + jj] = up[(b.HC - r) * WW + jj];  // every non-last CTA owns exactly HC rows
            if (dn) tmp[(b.nrows + 1 + r) * WW + jj] = r < down_rows ? dn[(1 + r) * WW + jj] : 0u;
        }
        cluster.sync();  // all pulls done before anyone overwrites tmp again
    } else {
        __syncthreads();
    }
    // Column pass.  An item is NRUN consecutive rows of one word column: its NRUN + 2K input words are loaded once
    // into registers and every output is K three-input ORs of them -- (NRUN + 2K) / NRUN shared-memory loads per
    // word instead of 2K + 1, which is what bounded this pass (one LDS per cycle and SM against four issue slots).
    // The last run of a column is anchored at the last row (it overlaps its neighbour and rewrites the same values).
    constexpr int NRUN = 10;
    if (b.nrows >= NRUN) {
        const int runs = (b.nrows + NRUN - 1) / NRUN;
        for (int item = tid; item < runs * WW; item += LNT) {
            const int q = item / WW, jj = item - q * WW;
            const int r0 = min(1 + q * NRUN, b.nrows - NRUN + 1);
            const uint32_t m = (jj == WW - 1) ? b.last_mask : 0xffffffffu;
            const uint32_t* p = tmp + (r0 - K) * WW + jj;
            uint32_t v[NRUN + 2 * K];
#pragma unroll
            for (int i = 0; i < NRUN + 2 * K; ++i) v[i] = p[i * WW];
            uint32_t* o = src + r0 * WW + jj;
#pragma unroll
            for (int i = 0; i < NRUN; ++i) {
                uint32_t acc = v[i];
#pragma unroll
                for (int d = 1; d <= K; ++d) acc |= v[i + 2 * d - 1] | v[i + 2 * d];
                o[i * WW] = (acc ^ inv) & m;
            }
        }
    } else if (active) {
        for (int lr = 1 + rb; lr <= b.nrows; lr += R) {
            const uint32_t* p = tmp + lr * WW + j;
            uint32_t acc = p[0];
#pragma unroll
            for (int d = 1; d <= K; ++d) acc |= p[-d * WW] | p[d * WW];
            src[lr * WW + j] = (acc ^ inv) & mC;
        }
    }
    __syncthreads();
}

template <int WWC>
__device__ __forceinline__ void morph_fast_k(cg::cluster_group& cluster, uint32_t* src, uint32_t* tmp, const Band& b, int k,
                                             bool erode, int CS)
{
    switch (k) {
    case 1: morph_fast_t<1, WWC>(cluster, src, tmp, b, erode, CS); break;
    case 2: morph_fast_t<2, WWC>(cluster, src, tmp, b, erode, CS); break;
    case 3: morph_fast_t<3, WWC>(cluster, src, tmp, b, erode, CS); break;
    case 4: morph_fast_t<4, WWC>(cluster, src, tmp, b, erode, CS); break;
    case 5: morph_fast_t<5, WWC>(cluster, src, tmp, b, erode, CS); break;
    default: morph_fast_t<8, WWC>(cluster, src, tmp, b, erode, CS); break;
    }
}

// k in {1, 2, 3, 4, 5, 8}: the iteration counts of the reference's chain (QuadDetection.h:213-214,
// HypothesisEvaluation.h:352-371)
__device__ __noinline__ void morph(cg::cluster_group& cluster, uint32_t* src, uint32_t* tmp, const Band& b, int k, bool erode,
                                   int CS)
{
    if (b.WW <= LNT / 2 && (CS == 1 || b.HC >= PADR)) {
        if (b.WW == 20) morph_fast_k<20>(cluster, src, tmp, b, k, erode, CS);
        else morph_fast_k<0>(cluster, src, tmp, b, k, erode, CS);
    } else {
        morph_op(cluster, src, tmp, b, k, erode, CS);
    }
}

__global__ void __launch_bounds__(LNT, 2) k_edge_link(LinkArgs a, int cluster_size)
{
    extern __shared__ uint32_t smem[];
    __shared__ int s_flags[2][8];

    cg::cluster_group cluster = cg::this_cluster();
    const int CS = cluster_size;
    Band b;
    b.W = a.W; b.H = a.H; b.WW = a.WW; b.HC = a.HC;
    b.rank = (CS > 1) ? (int)cluster.block_rank() : 0;
    b.row0 = b.rank * b.HC;
    b.nrows = max(0, min(b.HC, b.H - b.row0));
    b.last_rank = (b.H - 1) / b.HC;  // last CTA that owns rows
    b.last_mask = (b.W & 31) ? ((1u << (b.W & 31)) - 1u) : 0xffffffffu;
    const int frame = blockIdx.x / CS;
    const int WW = b.WW, nrows = b.nrows;

    // every plane: PADR - 1 zero rows, halo row 0, rows 1..HC, halo row HC + 1, PADR - 1 zero rows
    const int plane_words = (b.HC + 2 * PADR) * WW;
    uint32_t* A = smem + (PADR - 1) * WW;                 // weak, later ping-pong
    uint32_t* B = smem + plane_words + (PADR - 1) * WW;   // edges
    for (int i = threadIdx.x; i < (PADR - 1) * WW; i += LNT) {
        A[i - (PADR - 1) * WW] = 0u; B[i - (PADR - 1) * WW] = 0u;
        A[(b.HC + 2) * WW + i] = 0u; B[(b.HC + 2) * WW + i] = 0u;
    }

    const size_t plane = (size_t)frame * b.H * WW;
    for (int i = threadIdx.x; i < (b.HC + 2) * WW; i += LNT) {
        int r = i / WW - 1;  // local row index -1 .. HC
        uint32_t wv = 0, sv = 0;
        if (r >= 0 && r < nrows) {
            wv = a.weak[plane + (size_t)(b.row0 + r) * WW + (i - (r + 1) * WW)];
            sv = a.strong[plane + (size_t)(b.row0 + r) * WW + (i - (r + 1) * WW)];
        }
        A[i] = wv;
        B[i] = sv;
    }
    if (threadIdx.x < 16) (&s_flags[0][0])[threadIdx.x] = 0;
    __syncthreads();

    long long tk = clock64();
    auto tick = [&](int slot) {
        if (a.dbg && threadIdx.x == 0) { long long t = clock64(); atomicAdd((unsigned long long*)&a.dbg[slot], (unsigned long long)(t - tk)); tk = t; }
    };
    tick(4);
    // ---- hysteresis closure (in place, monotone: racing reads are benign) ----
    int parity = 0;
    for (;;) {
        for (;;) {
            int changed = 0;
            for (int i = threadIdx.x; i < nrows * WW; i += LNT) {
                int r = i / WW + 1, j = i - (r - 1) * WW;
                uint32_t m = A[r * WW + j];
                if (m == 0) continue;
                volatile uint32_t* p = B + r * WW + j;
                uint32_t c = p[0];
                if (c == m) continue;
                uint32_t nb = p[-WW] | c | p[WW];
                uint32_t l = (j > 0) ? (p[-WW - 1] | p[-1] | p[WW - 1]) >> 31 : 0u;
                uint32_t rr = (j < WW - 1) ? (p[-WW + 1] | p[1] | p[WW + 1]) << 31 : 0u;
                uint32_t seeds = (nb | (nb << 1) | (nb >> 1) | l | rr) & m;
                if (seeds == 0) continue;
                uint32_t f = flood_word(m, seeds);
                if (f != c) { p[0] = f; changed = 1; }
            }
            if (!__syncthreads_or(changed)) break;
        }
        if (CS == 1) break;
        cluster.sync();
        int hc = pull_halo(cluster, B, b);
        hc = __syncthreads_or(hc);
        if (threadIdx.x < CS) {
            int* remote = cluster.map_shared_rank(&s_flags[parity][b.rank], threadIdx.x);
            *remote = hc;
        }
        cluster.sync();
        int any = 0;
        for (int k = 0; k < CS; ++k) any |= s_flags[parity][k];
        parity ^= 1;
        if (!any) break;
    }

    tick(5);
    if (a.canny_u8) store_u8(B, b, a.canny_u8 + (size_t)frame * b.W * b.H);
    if (a.canny_bits) store_bits(B, b, a.canny_bits + plane);
    tick(8);

    const bool want_morph = a.morph_u8 || a.morph_bits;
    const bool want_mask = a.mask_u8 || a.mask_bits;
    if (!want_morph && !want_mask) return;  // uniform across the cluster

    if (want_morph) {
        // dilate x2, erode x1 on the closure result (B -> A -> B)
        if (CS > 1) cluster.sync(); else __syncthreads();
        morph(cluster, B, A, b, 2, false, CS);
        morph(cluster, B, A, b, 1, true, CS);
        tick(9);
        if (a.morph_u8) store_u8(B, b, a.morph_u8 + (size_t)frame * b.W * b.H);
        if (a.morph_bits) store_bits(B, b, a.morph_bits + plane);
    }
    tick(6);
    if (want_mask) {
        // the closure result was overwritten by the ping-pong: reload it (L2-resident) into B
        if (CS > 1) cluster.sync();  // neighbours may still be pulling from B / A
        __syncthreads();
        if (want_morph)
            for (int i = threadIdx.x; i < nrows * WW; i += LNT) B[WW + i] = a.canny_bits[plane + (size_t)b.row0 * WW + i];
        refresh_halo(cluster, B, b, CS);
        if (CS == 1 && WW <= LNT) {
            // One CTA holds the frame: a thread keeps one word column and strides over the rows, so that the
            // edge-of-image cases are loop-invariant masks (j) or one comparison (first / last row); the pad rows
            // and the words before / after a row exist in shared memory, so no load is guarded.
            const int R = LNT / WW, rb = threadIdx.x / WW, j = threadIdx.x - rb * WW;
            const bool act = rb < R;
            const uint32_t lm = (j == WW - 1) ? b.last_mask : 0xffffffffu;
            const uint32_t hasL = j > 0 ? 0xffffffffu : 0u, hasR = j + 1 < WW ? 0xffffffffu : 0u;
            const uint32_t padR = (j + 1 == WW - 1) ? ~b.last_mask : 0u;   // unused bits of the right neighbour word
            // t = NOT(cross-gradient(canny)) -> A   (dilation: 0 outside the image; erosion: 1 outside)
            if (act)
                for (int lr = 1 + rb; lr <= nrows; lr += R) {
                    const uint32_t* p = B + lr * WW + j;
                    const uint32_t u = p[-WW], c = p[0], d = p[WW], cl = p[-1], cr = p[1];
                    const uint32_t dil = u | c | d | ((c << 1) | ((cl & hasL) >> 31)) | ((c >> 1) | ((cr & hasR) << 31));
                    const uint32_t ce = c | ~lm, ue = lr == 1 ? 0xffffffffu : (u | ~lm), de = lr == nrows ? 0xffffffffu : (d | ~lm);
                    const uint32_t cle = cl | ~hasL, cre = (cr | padR) | ~hasR;
                    const uint32_t ero = ue & ce & de & ((ce << 1) | (cle >> 31)) & ((ce >> 1) | (cre << 31));
                    A[lr * WW + j] = ~(dil & ~ero) & lm;
                }
            __syncthreads();
            // mask0 = canny | (t & ~cross-interior(t)) -> B (in place over canny: reads A and own word only)
            if (act)
                for (int lr = 1 + rb; lr <= nrows; lr += R) {
                    const uint32_t* p = A + lr * WW + j;
                    const uint32_t t = p[0], u = lr == 1 ? 0u : p[-WW], d = lr == nrows ? 0u : p[WW];
                    const uint32_t interior = t & u & d & ((t << 1) | ((p[-1] & hasL) >> 31)) & ((t >> 1) | ((p[1] & hasR) << 31));
                    B[lr * WW + j] |= (t & ~interior);
                }
        } else {
            // t = NOT(cross-gradient(canny)) -> A
            for (int i = threadIdx.x; i < nrows * WW; i += LNT) {
                int lr = i / WW + 1, j = i - (lr - 1) * WW;
                uint32_t dil = stencil<true, false>(B, b, lr, j), ero = stencil<false, false>(B, b, lr, j);
                uint32_t t = ~(dil & ~ero);
                if (j == WW - 1) t &= b.last_mask;
                A[lr * WW + j] = t;
            }
            refresh_halo(cluster, A, b, CS);
            // mask0 = canny | (t & ~cross-interior(t)) -> B (in place over canny: reads A and own word only)
            for (int i = threadIdx.x; i < nrows * WW; i += LNT) {
                int lr = i / WW + 1, j = i - (lr - 1) * WW;
                uint32_t t = A[lr * WW + j];
                // interior = t & up & down & left & right with 0 outside the image
                uint32_t u = row_word(A, b, lr - 1, j, 0u), d = row_word(A, b, lr + 1, j, 0u);
                uint32_t cl = row_word(A, b, lr, j - 1, 0u), cr = row_word(A, b, lr, j + 1, 0u);
                uint32_t interior = t & u & d & ((t << 1) | (cl >> 31)) & ((t >> 1) | (cr << 31));
                B[lr * WW + j] |= (t & ~interior);
            }
        }
        // closes of 3, 4, 5 iterations then erode 3: dilate 3, erode 3, dilate 4, erode 4, dilate 5, erode 5 + 3
        if (CS > 1) cluster.sync(); else __syncthreads();
        morph(cluster, B, A, b, 3, false, CS);
        morph(cluster, B, A, b, 3, true, CS);
        morph(cluster, B, A, b, 4, false, CS);
        morph(cluster, B, A, b, 4, true, CS);
        morph(cluster, B, A, b, 5, false, CS);
        morph(cluster, B, A, b, 8, true, CS);
        if (a.mask_u8) store_u8(B, b, a.mask_u8 + (size_t)frame * b.W * b.H);
        if (a.mask_bits) store_bits(B, b, a.mask_bits + plane);
    }
    tick(7);
    // keep this CTA's shared memory alive until every neighbour has finished pulling from it
    if (CS > 1) cluster.sync();
}

}  // namespace

int launch_edge_link(mantis_ctx* c, const uint32_t* d_weak, const uint32_t* d_strong, int n, int w, int h,
                     const EdgeOutputs& out)
{
    const int WW = words_per_row(w);
    // smallest cluster whose per-CTA planes fit in ~100 KB (two CTAs per SM), else up to 200 KB
    int cs = 1;
    size_t smem = 0;
    const size_t want = 100 * 1024, hard = 200 * 1024;
    int chosen = -1;
    for (int pass = 0; pass < 2 && chosen < 0; ++pass)
        for (cs = 1; cs <= 8; cs *= 2) {
            int hc = (h + cs - 1) / cs;
            smem = (size_t)2 * (hc + 2 * PADR) * WW * sizeof(uint32_t);
            if (smem <= (pass == 0 ? want : hard)) { chosen = cs; break; }
        }
    if (chosen < 0) {
        snprintf(c->err, sizeof(c->err), "frame %dx%d too large for the on-chip edge stage", w, h);
        return MANTIS_ERR_UNSUPPORTED;
    }
    cs = chosen;
    LinkArgs a;
    a.weak = d_weak; a.strong = d_strong;
    a.canny_bits = out.canny_bits; a.morph_bits = out.morph_bits; a.mask_bits = out.mask_bits;
    a.canny_u8 = out.canny_u8; a.morph_u8 = out.morph_u8; a.mask_u8 = out.mask_u8;
    a.W = w; a.H = h; a.WW = WW; a.HC = (h + cs - 1) / cs;
    if ((a.mask_bits || a.mask_u8) && !a.canny_bits) return MANTIS_ERR_INVALID;
    a.dbg = (long long*)c->dbg_clocks;
    smem = (size_t)2 * (a.HC + 2 * PADR) * WW * sizeof(uint32_t);

    MCHECK(c, cudaFuncSetAttribute(k_edge_link, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(n * cs), 1, 1);
    cfg.blockDim = dim3(LNT, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = c->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    MCHECK(c, cudaLaunchKernelEx(&cfg, k_edge_link, a, cs));
    c->launches++;
    return MANTIS_OK;
}
