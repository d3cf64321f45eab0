// edge_candidates.cu -- K1: fused BGR->gray->blur->Sobel->L1 magnitude->NMS->double threshold.
//
// Realises, for a batch of frames, the arithmetic of
//   cv::cvtColor(BGR2GRAY)           include/mantis3/QuadDetection.h:209
//   cv::GaussianBlur(3x3, sigma=3)   :211
//   cv::Canny(T, 3T, 3) up to (not including) hysteresis   :212
// and emits two bit planes per frame (32 px per word, row-major):
//   weak   : pixel is a Canny candidate (mag > low and passes non-max suppression)
//   strong : candidate and mag > high
// The hysteresis closure and the morphology run on-chip in edge_link.cu.
//
// Integer spec (pinned against cv2 4.13 by the oracle, oracle/image_stage.c):
//   gray = (3735 B + 19235 G + 9798 R + 16384) >> 15
//   blur = taps (84,88,84)/256 per axis, BORDER_REFLECT_101, (v + 32768) >> 16
//   Sobel 3x3 on the blurred image with BORDER_REPLICATE; mag = |dx| + |dy|, 0 outside
//   NMS: TG22 = 13573 (Q15) sector test (phase 3 below).
//
// Shape of the kernel (one CTA of 4 warps per 128 x 32 output tile; HBM-bound by design: 3 B/px in,
// 2 bits/px out, every intermediate lives in shared memory as packed words of 4 px):
//   phase 0  BGR -> gray words.  Three 32-bit loads bring 4 px; each gray value is two IDP.4A dot
//            products (coefficients split as hi*128 + lo so that they fit a byte lane).
//   phase 1  gray -> blur.  Each lane owns one word column and marches down its warp's band of rows;
//            the horizontal taps are one IDP.4A per pixel on a funnel-shifted window, the vertical
//            taps run on a two-row register window.  (84,88,84) = 4 x (21,22,21): the factor 16 is
//            folded into the final shift, (21 s + 22 h + 2048) >> 12, exactly the same integer.
//   phase 2  blur -> (dx, dy) per pixel, stored as packed s16 pairs, same marching scheme; pixels
//            with |dx| + |dy| > low are appended to a compact candidate list.
//   phase 3  non-max suppression only for the listed pixels (a few percent of the tile), lanes =
//            candidates; results are OR-ed into the tile's two bit planes and written out.
#include "ctx.h"

namespace {

constexpr int TW = 128;           // output tile width (4 bit-plane words)
constexpr int TH = 32;            // output tile height
constexpr int NT = 128;           // threads per CTA (4 warps)
constexpr int GWW = TW / 4 + 2;   // gray / blur words per tile row: px x0-4 .. x0+TW+3
constexpr int GH = TH + 6;        // gray rows  y0-3 .. y0+TH+2
constexpr int BH = TH + 4;        // blur rows  y0-2 .. y0+TH+1
constexpr int DH = TH + 2;        // dxdy rows  y0-1 .. y0+TH
constexpr int DP = TW + 8;        // dxdy row pitch in px; px x0 sits at column 4 (16-byte aligned groups)

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
    return min(max(i, 0), n - 1);  // far outside (never used by in-image outputs)
}

__device__ __forceinline__ uint32_t gray_scalar(const uint8_t* p)
{
    return (uint32_t)((3735 * p[0] + 19235 * p[1] + 9798 * p[2] + 16384) >> 15);
}

// 2 x (3735 B + 19235 G + 9798 R + 16384) for the pixel in bytes 0..2 of `px` (byte 3 is ignored):
// the gray value is byte 2 of the result.
__device__ __forceinline__ uint32_t gray_x2(uint32_t px)
{
    // 3735 = 29*128 + 23, 19235 = 150*128 + 35, 9798 = 76*128 + 70; doubled low parts still fit a byte
    const uint32_t hi = __dp4a(px, 0x004C961Du, 0u);       // 29 B + 150 G + 76 R
    const uint32_t lo = __dp4a(px, 0x008C462Eu, 32768u);   // 46 B + 70 G + 140 R + 2*16384
    return hi * 256u + lo;
}

// horizontal (21,22,21) taps for the 4 px of word C; L / R are the neighbouring words
__device__ __forceinline__ void hrow_blur(uint32_t L, uint32_t C, uint32_t R, uint32_t h[4])
{
    const uint32_t K = 0x00151615u;
    h[0] = __dp4a(__funnelshift_r(L, C, 24), K, 0u);
    h[1] = __dp4a(C, K, 0u);
    h[2] = __dp4a(__funnelshift_r(C, R, 8), K, 0u);
    h[3] = __dp4a(__funnelshift_r(C, R, 16), K, 0u);
}

// unsigned bytes of a times signed bytes of b (IDP.4A.U8.S8)
__device__ __forceinline__ int dp4a_us(uint32_t a, int b)
{
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(0));
    return d;
}

// horizontal Sobel parts for the 4 px of word C: hs = l + 2c + r, hd = r - l
__device__ __forceinline__ void hrow_sobel(uint32_t L, uint32_t C, uint32_t R, int hs[4], int hd[4])
{
    const uint32_t w0 = __funnelshift_r(L, C, 24), w2 = __funnelshift_r(C, R, 8), w3 = __funnelshift_r(C, R, 16);
    const int KS = 0x00010201, KD = 0x000100FF;  // (1,2,1,0) and (-1,0,1,0) as signed bytes
    hs[0] = dp4a_us(w0, KS); hd[0] = dp4a_us(w0, KD);
    hs[1] = dp4a_us(C, KS);  hd[1] = dp4a_us(C, KD);
    hs[2] = dp4a_us(w2, KS); hd[2] = dp4a_us(w2, KD);
    hs[3] = dp4a_us(w3, KS); hd[3] = dp4a_us(w3, KD);
}

__device__ __forceinline__ uint32_t vblur_pack(const uint32_t a[4], const uint32_t b[4], const uint32_t c[4])
{
    // (21 (a + c) + 22 b + 2048) >> 12 per pixel, packed into one word
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = (21u * (a[k] + c[k]) + (22u * b[k] + 2048u)) >> 12;
    return v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24);
}

__device__ __forceinline__ int mag_of(uint32_t dxdy)
{
    const int dx = (int)(short)(dxdy & 0xffffu), dy = (int)dxdy >> 16;
    return abs(dx) + abs(dy);
}

struct Tile {
    uint32_t gray[GH][GWW];
    uint32_t blur[BH][GWW];
    uint32_t dxdy[DH][DP];
    uint16_t list[TW * TH];
    uint32_t weak[TH * (TW / 32)];
    uint32_t strong[TH * (TW / 32)];
    int count;
};

// ---- phase 2 body: Sobel on the blur tile.  Warp w marches over dxdy rows [9w, 9w+9) ∩ [0, DH) ----
template <bool BORDER>
__device__ __forceinline__ void sobel_march(Tile& s, int x0, int y0, int W, int H, int low, int warp, int lane)
{
    constexpr int RPW = 9;
    const int r0 = warp * RPW, lw = 1 + lane;
    int sa[4], da[4], sb[4], db[4], sc[4], dc[4];
    hrow_sobel(s.blur[r0][lw - 1], s.blur[r0][lw], s.blur[r0][lw + 1], sa, da);
    hrow_sobel(s.blur[r0 + 1][lw - 1], s.blur[r0 + 1][lw], s.blur[r0 + 1][lw + 1], sb, db);
    const int nr = min(RPW, DH - r0);  // 9, 9, 9, 7: warp-uniform
#pragma unroll 3
    for (int r = 0; r < nr; ++r) {
        const int ld = r0 + r;  // dxdy row ld <-> image row y0 - 1 + ld, needs blur rows ld .. ld+2
        hrow_sobel(s.blur[ld + 2][lw - 1], s.blur[ld + 2][lw], s.blur[ld + 2][lw + 1], sc, dc);
        int dx[4], dy[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            dx[k] = da[k] + 2 * db[k] + dc[k];
            dy[k] = sc[k] - sa[k];
        }
        if (BORDER) {
            const int gy = y0 - 1 + ld;
            const bool row_out = gy < 0 || gy >= H;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (row_out || x0 + 4 * lane + k >= W) { dx[k] = 0; dy[k] = 0; }
        }
        int m[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) m[k] = (int)__sad(dx[k], 0, __sad(dy[k], 0, 0u));
        *reinterpret_cast<uint4*>(&s.dxdy[ld][4 + 4 * lane]) =
            make_uint4(__byte_perm((uint32_t)dx[0], (uint32_t)dy[0], 0x5410), __byte_perm((uint32_t)dx[1], (uint32_t)dy[1], 0x5410),
                       __byte_perm((uint32_t)dx[2], (uint32_t)dy[2], 0x5410), __byte_perm((uint32_t)dx[3], (uint32_t)dy[3], 0x5410));
        // rows 0 and TH+1 are halo rows: no outputs there
        if (max(max(m[0], m[1]), max(m[2], m[3])) > low && ld >= 1 && ld <= TH) {
            uint32_t mask = (m[0] > low ? 1u : 0u) | (m[1] > low ? 2u : 0u) | (m[2] > low ? 4u : 0u) | (m[3] > low ? 8u : 0u);
            int o = atomicAdd(&s.count, __popc(mask));
            const int base = (ld - 1) * TW + 4 * lane;
            while (mask) {
                const int k = __ffs(mask) - 1;
                mask &= mask - 1;
                s.list[o++] = (uint16_t)(base + k);
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) { sa[k] = sb[k]; da[k] = db[k]; sb[k] = sc[k]; db[k] = dc[k]; }
    }
}

__global__ void __launch_bounds__(NT)
k_edge_candidates(const uint8_t* __restrict__ bgr, int W, int H, int stride, size_t frame_pitch,
                  int low, int high, uint32_t* __restrict__ weak, uint32_t* __restrict__ strong, int WW)
{
    __shared__ __align__(16) Tile s;

    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const uint8_t* img = bgr + (size_t)blockIdx.z * frame_pitch;
    const size_t plane = (size_t)blockIdx.z * H * WW;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // vector loads need 4-byte aligned rows and at least one whole 4-px group per row
    const bool vec_ok = ((((uintptr_t)img) | (uintptr_t)stride) & 3u) == 0 && W >= 4;
    // a tile whose whole halo lies inside the image needs no border handling at all
    const bool border = x0 < 4 || x0 + TW + 4 > W || y0 < 3 || y0 + TH + 3 > H;
    uint8_t* const gb = reinterpret_cast<uint8_t*>(&s.gray[0][0]);
    uint8_t* const bb = reinterpret_cast<uint8_t*>(&s.blur[0][0]);
    constexpr int GB = GWW * 4;  // bytes per gray / blur tile row

    if (tid == 0) s.count = 0;
    for (int i = tid; i < TH * (TW / 32); i += NT) { s.weak[i] = 0; s.strong[i] = 0; }

    // ---- phase 0: gray words ----
    if (vec_ok) {
        // Every word is computed from a legal, clamped address; words outside the image hold garbage
        // until the ring pass below fills the one-pixel frame that in-image outputs actually read.
        // Lane l owns word column 1 + l; warp w takes rows w, w+4, ...; loads are issued five rows at
        // a time so that each thread keeps 15 requests in flight.
        const int Wm4 = (W - 4) & ~3;
        auto gray_word = [](uint32_t w0, uint32_t w1, uint32_t w2) {
            const uint32_t t0 = gray_x2(w0), t1 = gray_x2(__funnelshift_r(w0, w1, 24));
            const uint32_t t2 = gray_x2(__funnelshift_r(w1, w2, 16)), t3 = gray_x2(w2 >> 8);
            return __byte_perm(__byte_perm(t0, t1, 0x0062), __byte_perm(t2, t3, 0x0062), 0x5410);
        };
        {
            const uint8_t* pcol = img + 3 * min(max(x0 + 4 * lane, 0), Wm4);
#pragma unroll 1
            for (int j0 = 0; j0 < 10; j0 += 5) {
                uint32_t w0[5], w1[5], w2[5];
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const int ly = min(warp + 4 * (j0 + k), GH - 1);
                    const int cy = min(max(y0 - 3 + ly, 0), H - 1);
                    const uint32_t* p = reinterpret_cast<const uint32_t*>(pcol + (size_t)cy * stride);
                    w0[k] = __ldg(p); w1[k] = __ldg(p + 1); w2[k] = __ldg(p + 2);
                }
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const int ly = warp + 4 * (j0 + k);
                    if (ly < GH) s.gray[ly][1 + lane] = gray_word(w0[k], w1[k], w2[k]);
                }
            }
        }
        if (tid < 2 * GH) {  // halo words 0 and GWW-1
            const int ly = tid >> 1, lw = (tid & 1) ? GWW - 1 : 0;
            const int cy = min(max(y0 - 3 + ly, 0), H - 1), cx = min(max(x0 - 4 + 4 * lw, 0), Wm4);
            const uint32_t* p = reinterpret_cast<const uint32_t*>(img + (size_t)cy * stride + 3 * cx);
            s.gray[ly][lw] = gray_word(__ldg(p), __ldg(p + 1), __ldg(p + 2));
        }
    } else {
        for (int i = tid; i < GH * GWW; i += NT) {
            const int ly = i / GWW, lw = i - ly * GWW;
            const uint8_t* row = img + (size_t)reflect101(y0 - 3 + ly, H) * stride;
            uint32_t g = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) g |= gray_scalar(row + 3 * reflect101(x0 - 4 + 4 * lw + k, W)) << (8 * k);
            s.gray[ly][lw] = g;
        }
    }
    __syncthreads();
    if (border && vec_ok) {
        if (W & 3) {  // the word straddling the right image edge: its in-image pixels, one thread per row
            const int gxs = W & ~3, lx = gxs - (x0 - 4);
            if (tid < GH && lx >= 0 && lx < GB) {
                const int gy = y0 - 3 + tid;
                if (gy >= 0 && gy < H)
                    for (int k = 0; gxs + k < W; ++k) gb[tid * GB + lx + k] = (uint8_t)gray_scalar(img + (size_t)gy * stride + 3 * (gxs + k));
            }
            __syncthreads();
        }
        // BORDER_REFLECT_101 ring: rows -1 and H over the in-image columns, columns -1 and W over rows -1..H
        for (int i = tid; i < 2 * GB + 2 * GH; i += NT) {
            int gy, gx;
            if (i < 2 * GB) {
                const int which = i >= GB;
                gx = x0 - 4 + (i - which * GB);
                gy = which ? H : -1;
                if (gx < 0 || gx >= W) continue;
            } else {
                const int j = i - 2 * GB, which = j >= GH;
                gy = y0 - 3 + (j - which * GH);
                gx = which ? W : -1;
                if (gy < -1 || gy > H) continue;
            }
            const int ly = gy - (y0 - 3), lx = gx - (x0 - 4);
            const int sy = reflect101(gy, H) - (y0 - 3), sx = reflect101(gx, W) - (x0 - 4);
            if (ly < 0 || ly >= GH || lx < 0 || lx >= GB || sy < 0 || sy >= GH || sx < 0 || sx >= GB) continue;
            gb[ly * GB + lx] = gb[sy * GB + sx];
        }
        __syncthreads();
    }

    // ---- phase 1: blur words.  Warp w marches over blur rows [9w, 9w+9); lane = word 1 + lane ----
    {
        constexpr int RPW = BH / 4;  // 9
        const int r0 = warp * RPW, lw = 1 + lane;
        uint32_t ha[4], hb[4], hc[4];
        hrow_blur(s.gray[r0][lw - 1], s.gray[r0][lw], s.gray[r0][lw + 1], ha);
        hrow_blur(s.gray[r0 + 1][lw - 1], s.gray[r0 + 1][lw], s.gray[r0 + 1][lw + 1], hb);
#pragma unroll 3
        for (int r = 0; r < RPW; ++r) {
            const int gr = r0 + r + 2;
            hrow_blur(s.gray[gr][lw - 1], s.gray[gr][lw], s.gray[gr][lw + 1], hc);
            s.blur[r0 + r][lw] = vblur_pack(ha, hb, hc);
#pragma unroll
            for (int k = 0; k < 4; ++k) { ha[k] = hb[k]; hb[k] = hc[k]; }
        }
        // halo words 0 and GWW-1 (only their two inner pixels are consumed): one thread per (row, side)
        if (tid < 2 * BH) {
            const int br = tid >> 1, hw = (tid & 1) ? GWW - 1 : 0;
            uint32_t a[4], b[4], c[4];
            const uint32_t zl = 0u;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                uint32_t* dst = k == 0 ? a : (k == 1 ? b : c);
                const uint32_t L = hw ? s.gray[br + k][hw - 1] : zl, R = hw ? zl : s.gray[br + k][hw + 1];
                hrow_blur(L, s.gray[br + k][hw], R, dst);
            }
            s.blur[br][hw] = vblur_pack(a, b, c);
        }
    }
    __syncthreads();

    // BORDER_REPLICATE for Sobel: blurred values outside the image are those at the clamped coordinate.
    // Columns first (in-image rows), then whole rows copied from the clamped row.
    if (border) {
        const int lx0 = max(0, -(x0 - 4));                  // byte column of image x = 0 (if the tile reaches it)
        const int lxR = min(GB - 1, W - 1 - (x0 - 4));      // byte column of image x = W-1
        const int lbT = max(0, -(y0 - 2));                  // first in-image blur row
        const int lbB = min(BH - 1, H - 1 - (y0 - 2));      // last in-image blur row
        if (tid >= lbT && tid <= lbB) {
            uint8_t* row = bb + tid * GB;
            for (int c = 0; c < lx0; ++c) row[c] = row[lx0];
            for (int c = lxR + 1; c < GB; ++c) row[c] = row[lxR];
        }
        __syncthreads();
        const int n_out = lbT + (BH - 1 - lbB);
        for (int i = tid; i < n_out * GWW; i += NT) {
            int r = i / GWW;
            const int w = i - r * GWW;
            if (r >= lbT) r = lbB + 1 + (r - lbT);
            s.blur[r][w] = s.blur[r < lbT ? lbT : lbB][w];
        }
        __syncthreads();
    }

    // ---- phase 2: Sobel (dx, dy) + candidate list ----
    if (border) sobel_march<true>(s, x0, y0, W, H, low, warp, lane);
    else sobel_march<false>(s, x0, y0, W, H, low, warp, lane);
    // halo columns x0-1 and x0+TW: one thread per (row, side), straight 3x3 on blur bytes
    if (tid < 2 * DH) {
        const int ld = tid >> 1, side = tid & 1;
        const int gx = side ? x0 + TW : x0 - 1, gy = y0 - 1 + ld;
        uint32_t v = 0;
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            const int cx = gx - (x0 - 4);  // byte column in the blur tile
            const uint8_t* r0p = bb + ld * GB + cx;  // blur row ld = image row gy - 1
            const uint8_t* r1p = r0p + GB;
            const uint8_t* r2p = r1p + GB;
            const int a = r0p[-1], b = r0p[0], c = r0p[1], d = r1p[-1], f = r1p[1], g = r2p[-1], hh = r2p[0],
                      kk = r2p[1];
            const int dx = (c - a) + 2 * (f - d) + (kk - g);
            const int dy = (g + 2 * hh + kk) - (a + 2 * b + c);
            v = __byte_perm((uint32_t)dx, (uint32_t)dy, 0x5410);
        }
        s.dxdy[ld][side ? 4 + TW : 3] = v;
    }
    __syncthreads();

    // ---- phase 3: non-max suppression + double threshold for the listed pixels ----
    const int n = s.count;
    for (int i = tid; i < n; i += NT) {
        const int pos = s.list[i];
        const int ty = pos >> 7, tx = pos & (TW - 1);
        const uint32_t* c = &s.dxdy[ty + 1][tx + 4];
        const uint32_t w = c[0];
        const int xs = (int)(short)(w & 0xffffu), ys = (int)w >> 16;
        const int ax = abs(xs), ay = abs(ys) << 15;
        const int m = abs(xs) + abs(ys);
        const int tg22x = ax * 13573;
        // sector -> the two neighbours to beat: horizontal (left strictly, right or equal), vertical (up
        // strictly, down or equal), diagonal (both strictly); one pair of loads whatever the sector
        const bool horiz = ay < tg22x, vert = ay > tg22x + (ax << 16);
        const int sgn = ((xs ^ ys) < 0) ? -1 : 1;
        const int o1 = horiz ? -1 : (vert ? -DP : -DP - sgn);
        const int m1 = mag_of(c[o1]), m2 = mag_of(c[-o1]);
        const bool cand = (m > m1) && ((horiz || vert) ? (m >= m2) : (m > m2));
        if (cand) {
            const int wi = ty * (TW / 32) + (tx >> 5);
            const uint32_t bit = 1u << (tx & 31);
            atomicOr(&s.weak[wi], bit);
            if (m > high) atomicOr(&s.strong[wi], bit);
        }
    }
    __syncthreads();

    for (int i = tid; i < TH * (TW / 32); i += NT) {
        const int ty = i >> 2, gy = y0 + ty, wj = (x0 >> 5) + (i & 3);
        if (gy < H && wj < WW) {
            weak[plane + (size_t)gy * WW + wj] = s.weak[i];
            strong[plane + (size_t)gy * WW + wj] = s.strong[i];
        }
    }
}

}  // namespace

int launch_edge_candidates(mantis_ctx* c, const uint8_t* d_bgr, int n, int w, int h, int stride,
                           size_t frame_pitch, int low, int high, uint32_t* d_weak, uint32_t* d_strong)
{
    dim3 grid((w + TW - 1) / TW, (h + TH - 1) / TH, n);
    k_edge_candidates<<<grid, NT, 0, c->stream>>>(d_bgr, w, h, stride, frame_pitch, low, high, d_weak,
                                                  d_strong, words_per_row(w));
    c->launches++;
    MCHECK(c, cudaGetLastError());
    return MANTIS_OK;
}
