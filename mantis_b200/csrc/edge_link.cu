// edge_link.cu -- K2: per-frame on-chip Canny hysteresis closure + dilate x2 + erode x1.
//
// Completes cv::Canny (the 8-connected hysteresis closure, QuadDetection.h:212) and applies
// cv::dilate(iter 2) / cv::erode(iter 1) (:213-214) on bit-packed planes that never leave
// shared memory: one thread-block cluster per frame, the rows of the frame split across
// the CTAs of the cluster, halo rows exchanged through distributed shared memory.
//
//   closure : E <- strong;  repeat E <- weak & flood(E | dilate3x3(E)) until stable.  In-word
//             horizontal runs are filled in O(1) with the carry trick ((M+S)^M)&M.
//   dilate  : 3x3 max, out-of-image pixels never win  == OR with zero halo
//   erode   : 3x3 min, out-of-image pixels never win  == AND with all-ones halo
#include <cooperative_groups.h>
#include "ctx.h"

namespace cg = cooperative_groups;

namespace {

constexpr int LNT = 512;

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
    uint32_t* canny_bits;
    uint32_t* morph_bits;
    uint8_t* canny_u8;
    uint8_t* morph_u8;
    int W, H, WW, HC;  // HC = rows per CTA
};

// write rows [row0, row0+nrows) of a bit plane held in smem (local rows 1..nrows) as u8 {0,255}
__device__ void store_u8(const uint32_t* s, int WW, int W, int row0, int nrows, uint8_t* dst)
{
    if ((W & 15) == 0 && (((uintptr_t)dst) & 15) == 0) {
        int halves = W >> 4;  // 16-px groups per row
        for (int i = threadIdx.x; i < nrows * halves; i += LNT) {
            int r = i / halves, hx = i - r * halves;
            uint32_t w = s[(r + 1) * WW + (hx >> 1)];
            uint32_t b = (hx & 1) ? (w >> 16) : (w & 0xffffu);
            uint4 v;
            v.x = expand4(b & 15u);
            v.y = expand4((b >> 4) & 15u);
            v.z = expand4((b >> 8) & 15u);
            v.w = expand4((b >> 12) & 15u);
            *reinterpret_cast<uint4*>(dst + (size_t)(row0 + r) * W + hx * 16) = v;
        }
    } else {
        for (int i = threadIdx.x; i < nrows * W; i += LNT) {
            int r = i / W, x = i - r * W;
            uint32_t w = s[(r + 1) * WW + (x >> 5)];
            dst[(size_t)(row0 + r) * W + x] = ((w >> (x & 31)) & 1u) ? 255 : 0;
        }
    }
}

// Pull the neighbours' boundary rows of plane `s` (same smem offset in every CTA) into my halo
// rows 0 and nrows+1.  `fill` is the value used where there is no neighbour (image border).
// Returns (per thread) whether a halo word changed.
__device__ int pull_halo(cg::cluster_group& cluster, uint32_t* s, int WW, int HC, int nrows, int rank,
                         int last_rank, const int* s_nrows_dummy, uint32_t fill)
{
    (void)s_nrows_dummy;
    int changed = 0;
    for (int j = threadIdx.x; j < 2 * WW; j += LNT) {
        int top = j < WW;
        int jj = top ? j : j - WW;
        uint32_t v = fill;
        if (top) {
            if (rank > 0) {
                const uint32_t* nb = cluster.map_shared_rank(s, rank - 1);
                v = nb[HC * WW + jj];  // every non-last CTA owns exactly HC rows
            }
        } else {
            if (rank < last_rank) {
                const uint32_t* nb = cluster.map_shared_rank(s, rank + 1);
                v = nb[1 * WW + jj];
            }
        }
        uint32_t* dst = top ? &s[jj] : &s[(nrows + 1) * WW + jj];
        if (*dst != v) { *dst = v; changed = 1; }
    }
    return changed;
}

template <bool DILATE>
__device__ void morph3(const uint32_t* src, uint32_t* dst, int WW, int nrows, uint32_t last_mask)
{
    for (int i = threadIdx.x; i < nrows * WW; i += LNT) {
        int r = i / WW + 1, j = i - (r - 1) * WW;
        const uint32_t* p = src + r * WW + j;
        uint32_t vc, vl, vr;
        if (DILATE) {
            vc = p[-WW] | p[0] | p[WW];
            vl = (j > 0) ? (p[-WW - 1] | p[-1] | p[WW - 1]) >> 31 : 0u;
            vr = (j < WW - 1) ? (p[-WW + 1] | p[1] | p[WW + 1]) << 31 : 0u;
            uint32_t o = vc | (vc << 1) | (vc >> 1) | vl | vr;
            if (j == WW - 1) o &= last_mask;
            dst[r * WW + j] = o;
        } else {
            uint32_t inv = (j == WW - 1) ? ~last_mask : 0u;  // out-of-image columns never win the min
            vc = (p[-WW] | inv) & (p[0] | inv) & (p[WW] | inv);
            vl = (j > 0) ? (p[-WW - 1] & p[-1] & p[WW - 1]) >> 31 : 1u;
            uint32_t invr = (j + 1 == WW - 1) ? ~last_mask : 0u;
            vr = (j < WW - 1) ? ((p[-WW + 1] | invr) & (p[1] | invr) & (p[WW + 1] | invr)) << 31 : 0x80000000u;
            uint32_t o = vc & ((vc << 1) | vl) & ((vc >> 1) | vr);
            if (j == WW - 1) o &= last_mask;
            dst[r * WW + j] = o;
        }
    }
}

__global__ void __launch_bounds__(LNT) k_edge_link(LinkArgs a, int cluster_size)
{
    extern __shared__ uint32_t smem[];
    __shared__ int s_flags[2][8];

    cg::cluster_group cluster = cg::this_cluster();
    const int CS = cluster_size;
    const int rank = (CS > 1) ? (int)cluster.block_rank() : 0;
    const int frame = blockIdx.x / CS;
    const int W = a.W, H = a.H, WW = a.WW, HC = a.HC;
    const int row0 = rank * HC;
    const int nrows = max(0, min(HC, H - row0));
    // last CTA that owns rows
    const int last_rank = (H - 1) / HC;
    const uint32_t last_mask = (W & 31) ? ((1u << (W & 31)) - 1u) : 0xffffffffu;

    uint32_t* A = smem;                    // weak, later morphology ping-pong
    uint32_t* B = smem + (HC + 2) * WW;    // edges

    const size_t plane = (size_t)frame * H * WW;
    for (int i = threadIdx.x; i < (HC + 2) * WW; i += LNT) {
        int r = i / WW - 1;  // local row index -1 .. HC
        uint32_t wv = 0, sv = 0;
        if (r >= 0 && r < nrows) {
            wv = a.weak[plane + (size_t)(row0 + r) * WW + (i - (r + 1) * WW)];
            sv = a.strong[plane + (size_t)(row0 + r) * WW + (i - (r + 1) * WW)];
        }
        A[i] = wv;
        B[i] = sv;
    }
    if (threadIdx.x < 16) (&s_flags[0][0])[threadIdx.x] = 0;
    __syncthreads();

    // ---- hysteresis closure ----
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
        int hc = pull_halo(cluster, B, WW, HC, nrows, rank, last_rank, nullptr, 0u);
        hc = __syncthreads_or(hc);
        if (threadIdx.x < CS) {
            int* remote = cluster.map_shared_rank(&s_flags[parity][rank], threadIdx.x);
            *remote = hc;
        }
        cluster.sync();
        int any = 0;
        for (int k = 0; k < CS; ++k) any |= s_flags[parity][k];
        parity ^= 1;
        if (!any) break;
    }

    if (a.canny_u8) store_u8(B, WW, W, row0, nrows, a.canny_u8 + (size_t)frame * W * H);
    if (a.canny_bits)
        for (int i = threadIdx.x; i < nrows * WW; i += LNT)
            a.canny_bits[plane + (size_t)row0 * WW + i] = B[WW + i];

    if (a.morph_u8 == nullptr && a.morph_bits == nullptr) return;  // uniform across the cluster

    // ---- dilate x2, erode x1 (ping-pong B -> A -> B -> A) ----
    // halo rows of B are current (closure left them exchanged); image-border halos are 0.
    __syncthreads();
    morph3<true>(B, A, WW, nrows, last_mask);
    if (CS > 1) { cluster.sync(); pull_halo(cluster, A, WW, HC, nrows, rank, last_rank, nullptr, 0u); }
    else {
        __syncthreads();
        for (int j = threadIdx.x; j < WW; j += LNT) { A[j] = 0u; A[(nrows + 1) * WW + j] = 0u; }
    }
    __syncthreads();
    morph3<true>(A, B, WW, nrows, last_mask);
    if (CS > 1) { cluster.sync(); pull_halo(cluster, B, WW, HC, nrows, rank, last_rank, nullptr, 0xffffffffu); }
    else {
        __syncthreads();
        for (int j = threadIdx.x; j < WW; j += LNT) { B[j] = 0xffffffffu; B[(nrows + 1) * WW + j] = 0xffffffffu; }
    }
    __syncthreads();
    morph3<false>(B, A, WW, nrows, last_mask);
    __syncthreads();

    if (a.morph_u8) store_u8(A, WW, W, row0, nrows, a.morph_u8 + (size_t)frame * W * H);
    if (a.morph_bits)
        for (int i = threadIdx.x; i < nrows * WW; i += LNT)
            a.morph_bits[plane + (size_t)row0 * WW + i] = A[WW + i];
    // keep this CTA's shared memory alive until every neighbour has finished pulling from it
    if (CS > 1) cluster.sync();
}

}  // namespace

int launch_edge_link(mantis_ctx* c, const uint32_t* d_weak, const uint32_t* d_strong, int n, int w, int h,
                     uint32_t* d_canny_bits, uint32_t* d_morph_bits, uint8_t* d_canny_u8, uint8_t* d_morph_u8)
{
    const int WW = words_per_row(w);
    // smallest cluster whose per-CTA planes fit in ~100 KB (two CTAs per SM), else up to 227 KB
    int cs = 1;
    size_t smem = 0;
    const size_t want = 100 * 1024, hard = 200 * 1024;
    int chosen = -1;
    for (int pass = 0; pass < 2 && chosen < 0; ++pass)
        for (cs = 1; cs <= 8; cs *= 2) {
            int hc = (h + cs - 1) / cs;
            smem = (size_t)2 * (hc + 2) * WW * sizeof(uint32_t);
            if (smem <= (pass == 0 ? want : hard)) { chosen = cs; break; }
        }
    if (chosen < 0) {
        snprintf(c->err, sizeof(c->err), "frame %dx%d too large for the on-chip edge stage", w, h);
        return MANTIS_ERR_UNSUPPORTED;
    }
    cs = chosen;
    LinkArgs a;
    a.weak = d_weak; a.strong = d_strong;
    a.canny_bits = d_canny_bits; a.morph_bits = d_morph_bits;
    a.canny_u8 = d_canny_u8; a.morph_u8 = d_morph_u8;
    a.W = w; a.H = h; a.WW = WW; a.HC = (h + cs - 1) / cs;
    smem = (size_t)2 * (a.HC + 2) * WW * sizeof(uint32_t);

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
