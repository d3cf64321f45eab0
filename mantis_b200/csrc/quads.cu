// quads.cu -- K5: grid-cell (quad) extraction from the bit-packed morphology image.
//
// Realises, per frame and with OpenCV's exact results AND output order,
//   cv::findContours(RETR_CCOMP, CHAIN_APPROX_SIMPLE)       include/mantis3/QuadDetection.h:216
//   cv::approxPolyDP(contour, POLYGON_EPSILON, closed)      :221   (keep 4-gons, :222-225)
//   Quadrilateral ctor (centre, side, stretched test points)  :13-66
//   removeDuplicateQuads                                     :115-171
//
// Suzuki-Abe border following is a sequential raster scan in OpenCV.  Here every border of every
// frame of the batch is found and traced independently, in five flat launches:
//   k_qcand   word-parallel candidate detection.  A border can only start at the raster-first pixel
//             of its component: an outer border at the first pixel of a foreground run that has no
//             foreground anywhere above it (8-connectivity: columns start-1 .. end+1 of the row
//             above), a hole border at the first pixel of a background run whose left neighbour is
//             foreground and that has no background above it (4-connectivity: columns start .. end).
//   k_qprobe  one thread per candidate follows its border and gives up as soon as it meets a pixel
//             that the raster scan would have reached earlier (or, for holes, the image frame); a
//             candidate that closes its loop is exactly a border start of the sequential algorithm.
//   k_qrank   per frame: starts ranked in raster order (= OpenCV's discovery order), point offsets,
//             and a by-length order so that the warps of the next launch get similar borders.
//   k_qtrace  one thread per border: second walk emitting the CHAIN_APPROX_SIMPLE vertices and
//             recording which border owns each left edge it passes (hole -> component relation of
//             RETR_CCOMP), then approxPolyDP (OpenCV 4.13 variant: distance to the SEGMENT).
//   k_qfinish per frame: hole parents, CCOMP ordering (outer borders in reverse discovery order,
//             each followed by its holes, reversed), Quadrilateral ctor, de-duplication, output.
// That is the fallback.  Frames whose bit plane fits in shared memory take the segment-parallel
// route (k_qcut, k_qseg, k_qlink, k_qemit below): borders are cut into short pieces that are walked
// independently and linked afterwards, so no thread ever follows a whole border.
// A walk step is branch-light: the 3x3 neighbourhood comes from three word loads, the next
// direction from a rotate + find-first-set on the 8-bit neighbour mask.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ctx.h"
#include "pipeline.cuh"

namespace {

constexpr int MAXB = MANTIS_MAX_BORDERS;   // border starts per frame
constexpr int MAXCAND = 16384;             // candidates per frame
constexpr int CNT = 256;

struct BitImg {
    const uint32_t* bits;
    int W, H, WW;
    // bits x-1, x, x+1 of row y (0 outside the image) in bits 0..2
    __device__ __forceinline__ uint32_t get3(int x, int y) const
    {
        if ((unsigned)y >= (unsigned)H) return 0u;
        const uint32_t* row = bits + (size_t)y * WW;
        const int j = x >> 5, b = x & 31;
        uint32_t w = row[j], v;
        if (b == 0) v = (w << 1) | (j > 0 ? (row[j - 1] >> 31) : 0u);
        else if (b == 31) v = (w >> 30) | ((j + 1 < WW ? (row[j + 1] & 1u) : 0u) << 2);
        else v = w >> (b - 1);
        return v & 7u;
    }
    // 8-neighbour mask of (x, y): bit d set iff the pixel in direction d is foreground
    // directions: 0 E, 1 NE, 2 N, 3 NW, 4 W, 5 SW, 6 S, 7 SE (image coordinates, y down)
    __device__ __forceinline__ uint32_t nbr(int x, int y) const
    {
        uint32_t up = get3(x, y - 1), mid = get3(x, y), dn = get3(x, y + 1);
        return ((mid >> 2) & 1u) | (((up >> 2) & 1u) << 1) | (((up >> 1) & 1u) << 2) | ((up & 1u) << 3) |
               ((mid & 1u) << 4) | ((dn & 1u) << 5) | (((dn >> 1) & 1u) << 6) | (((dn >> 2) & 1u) << 7);
    }
    __device__ __forceinline__ int at(int x, int y) const
    {
        if ((unsigned)x >= (unsigned)W || (unsigned)y >= (unsigned)H) return 0;
        return (bits[(size_t)y * WW + (x >> 5)] >> (x & 31)) & 1;
    }
};

// The same plane staged in shared memory with a zero frame: one zero row above and below, one
// zero word left and right of every row (stride = WW + 2 words).  No bounds checks, no branches:
// the three neighbourhood rows are funnel shifts of two adjacent words each.
struct PadImg {
    const uint32_t* base;  // word (row -1, word -1) of the padded plane
    int stride;
    __device__ __forceinline__ uint32_t nbr(int x, int y) const
    {
        const int xm = x + 31;  // column x-1 in padded coordinates
        const uint32_t* p = base + (y + 1) * stride + (xm >> 5);
        const int b = xm & 31;
        const uint32_t up = __funnelshift_r(p[-stride], p[-stride + 1], b) & 7u;
        const uint32_t mid = __funnelshift_r(p[0], p[1], b) & 7u;
        const uint32_t dn = __funnelshift_r(p[stride], p[stride + 1], b) & 7u;
        // up: NW N NE, mid: W . E, dn: SW S SE  ->  direction order E NE N NW W SW S SE
        return (mid >> 2) | ((up & 4u) >> 1) | ((up & 2u) << 1) | ((up & 1u) << 3) | ((mid & 1u) << 4) | (dn << 5);
    }
    __device__ __forceinline__ uint32_t at(int x, int y) const  // x in [-32, W + 31], y in [-1, H]
    {
        return (base[(y + 1) * stride + ((x + 32) >> 5)] >> ((x + 32) & 31)) & 1u;
    }
};

__device__ __forceinline__ int dir_dx(int s) { return (int)((0x901Au >> (2 * s)) & 3u) - 1; }
__device__ __forceinline__ int dir_dy(int s) { return (int)((0xA901u >> (2 * s)) & 3u) - 1; }

// Border following (OpenCV's icvFetchContour on a 0/1 image with a zero frame).
// Visitor:  bool step(x, y, zmask)  -- called for every border pixel with the set of directions
//           examined as background from it; return false to abort.
//           void point(x, y)        -- CHAIN_APPROX_SIMPLE vertex.
// Returns true if the loop closed.
template <class Img, class V>
__device__ __forceinline__ bool follow_border(const Img& im, int x0, int y0, bool is_hole, V& vis)
{
    const int s_end = is_hole ? 0 : 4;
    uint32_t nb = im.nbr(x0, y0);
    int s = -1;
    for (int k = 1; k < 8; ++k) {  // clockwise from s_end - 1
        int d = (s_end - k) & 7;
        if ((nb >> d) & 1u) { s = d; break; }
    }
    if (s < 0) {  // isolated pixel
        if (!vis.step(x0, y0, 0xffu)) return false;
        vis.point(x0, y0);
        return true;
    }
    const int x1 = x0 + dir_dx(s), y1 = y0 + dir_dy(s);
    int x3 = x0, y3 = y0, prev_s = s ^ 4;
    for (;;) {
        // counter-clockwise from s + 1: first foreground neighbour
        const int r0 = (s + 1) & 7;
        const uint32_t rot = ((nb | (nb << 8)) >> r0) & 0xffu;
        const int k = __ffs(rot) - 1;
        uint32_t z = ((1u << k) - 1u) << r0;
        z = (z | (z >> 8)) & 0xffu;
        s = (r0 + k) & 7;
        if (!vis.step(x3, y3, z)) return false;
        if (s != prev_s) {
            vis.point(x3, y3);
            prev_s = s;
        }
        const int x4 = x3 + dir_dx(s), y4 = y3 + dir_dy(s);
        if (x4 == x0 && y4 == y0 && x3 == x1 && y3 == y1) break;
        x3 = x4;
        y3 = y4;
        s = (s + 4) & 7;
        nb = im.nbr(x3, y3);
    }
    return true;
}

// detection visitor: outer borders
struct OuterProbe {
    int x0, y0, npts, steps;
    __device__ __forceinline__ bool step(int x, int y, unsigned)
    {
        ++steps;
        return !(y < y0 || (y == y0 && x < x0));
    }
    __device__ __forceinline__ void point(int, int) { ++npts; }
};
// detection visitor: hole borders; (hx, hy) is the hole's first background pixel
struct HoleProbe {
    int hx, hy, W, H, npts, steps;
    __device__ __forceinline__ bool step(int x, int y, unsigned zmask)
    {
        ++steps;
#pragma unroll
        for (int d = 0; d < 8; d += 2) {
            if (!(zmask & (1u << d))) continue;
            int bx = x + dir_dx(d), by = y + dir_dy(d);
            if ((unsigned)bx >= (unsigned)W || (unsigned)by >= (unsigned)H) return false;  // touches the frame
            if (by < hy || (by == hy && bx < hx)) return false;
        }
        return true;
    }
    __device__ __forceinline__ void point(int, int) { ++npts; }
};
// extraction visitor
struct Emit {
    short2* out;
    uint16_t* owner;
    int W, n;
    uint16_t id;
    __device__ __forceinline__ bool step(int x, int y, unsigned zmask)
    {
        if (zmask & (1u << 4)) owner[(size_t)y * W + x] = id;  // this border owns the left edge of (x, y)
        return true;
    }
    __device__ __forceinline__ void point(int x, int y) { out[n++] = make_short2((short)x, (short)y); }
};
// approxPolyDP's final pass: drops (almost) collinear vertices of dst[0..count), in place.
__device__ int approx_cleanup(short2* dst, int count, double eps)
{
    int new_count = count;
    short2 start_pt, end_pt, pt;
    int pos = count - 1;
    start_pt = dst[pos];
    if (++pos >= count) pos = 0;
    int wpos = pos;
    pt = dst[pos];
    if (++pos >= count) pos = 0;
    for (int i = 0; i < count && new_count > 2; ++i) {
        end_pt = dst[pos];
        if (++pos >= count) pos = 0;
        double dx = end_pt.x - start_pt.x, dy = end_pt.y - start_pt.y;
        double dist = fabs((pt.x - start_pt.x) * dy - (pt.y - start_pt.y) * dx);
        double sip = (double)(pt.x - start_pt.x) * (end_pt.x - pt.x) + (double)(pt.y - start_pt.y) * (end_pt.y - pt.y);
        if (dist * dist <= 0.5 * eps * (dx * dx + dy * dy) && dx != 0 && dy != 0 && sip >= 0) {
            new_count--;
            dst[wpos] = start_pt = end_pt;
            if (++wpos >= count) wpos = 0;
            pt = dst[pos];
            if (++pos >= count) pos = 0;
            i++;
            continue;
        }
        dst[wpos] = start_pt = pt;
        if (++wpos >= count) wpos = 0;
        pt = end_pt;
    }
    return new_count;
}

// cv::approxPolyDP for a closed int contour (see oracle/approx.c for the algorithm notes).
__device__ int approx_poly_closed(const short2* src, int count, double eps, short2* dst, int2* stack)
{
    if (count == 0) return 0;
    int new_count = 0, top = 0, pos = 0;
    int2 slice = make_int2(0, 0), right = make_int2(0, 0);
    short2 start_pt = make_short2(0, 0), end_pt, pt;
    bool le_eps = false;
    eps *= eps;
    right.x = 0;
    for (int i = 0; i < 3; ++i) {
        double max_dist = 0;
        pos = (pos + right.x) % count;
        start_pt = src[pos];
        if (++pos >= count) pos = 0;
        for (int j = 1; j < count; ++j) {
            pt = src[pos];
            if (++pos >= count) pos = 0;
            double dx = pt.x - start_pt.x, dy = pt.y - start_pt.y;
            double dist = dx * dx + dy * dy;
            if (dist > max_dist) { max_dist = dist; right.x = j; }
        }
        le_eps = max_dist <= eps;
    }
    if (!le_eps) {
        right.y = slice.x = pos % count;
        slice.y = right.x = (right.x + slice.x) % count;
        stack[top++] = right;
        stack[top++] = slice;
    } else {
        dst[new_count++] = start_pt;
    }
    while (top > 0) {
        slice = stack[--top];
        end_pt = src[slice.y];
        pos = slice.x;
        start_pt = src[pos];
        if (++pos >= count) pos = 0;
        if (pos != slice.y) {
            double dx = end_pt.x - start_pt.x, dy = end_pt.y - start_pt.y;
            double L = dx * dx + dy * dy, max_dist = 0;
            while (pos != slice.y) {
                pt = src[pos];
                if (++pos >= count) pos = 0;
                double px = pt.x - start_pt.x, py = pt.y - start_pt.y;
                double t = px * dx + py * dy, dist;
                if (t <= 0) dist = (px * px + py * py) * L;
                else if (t >= L) { double qx = pt.x - end_pt.x, qy = pt.y - end_pt.y; dist = (qx * qx + qy * qy) * L; }
                else { double cr = py * dx - px * dy; dist = cr * cr; }
                if (dist > max_dist) { max_dist = dist; right.x = (pos + count - 1) % count; }
            }
            le_eps = max_dist <= eps * L;
        } else {
            le_eps = true;
            start_pt = src[slice.x];
        }
        if (le_eps) {
            dst[new_count++] = start_pt;
        } else {
            right.y = slice.y;
            slice.y = right.x;
            stack[top++] = right;
            stack[top++] = slice;
        }
    }
    return approx_cleanup(dst, new_count, eps);
}


// (dist, order) arg-max over the warp: larger dist wins, ties go to the smaller order (= the
// element a sequential `dist > max_dist` scan would have kept).
__device__ __forceinline__ void warp_argmax_first(double& d, int& ord)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        double d2 = __shfl_xor_sync(0xffffffffu, d, o);
        int o2 = __shfl_xor_sync(0xffffffffu, ord, o);
        if (d2 > d || (d2 == d && o2 < ord)) { d = d2; ord = o2; }
    }
}

// approx_poly_closed with the point scans spread over a warp (all 32 lanes call it with the same
// arguments; the result is identical to the sequential version).  Used for long contours.
__device__ int approx_poly_closed_warp(const short2* src, int count, double eps, short2* dst, int2* stack)
{
    if (count == 0) return 0;
    const int lane = threadIdx.x & 31;
    int new_count = 0, top = 0, pos = 0;
    int2 slice = make_int2(0, 0), right = make_int2(0, 0);
    short2 start_pt = make_short2(0, 0), end_pt;
    bool le_eps = false;
    eps *= eps;
    for (int i = 0; i < 3; ++i) {
        pos = (pos + right.x) % count;     // index of the start point; the scan returns to it
        start_pt = src[pos];
        double best = 0;
        int bj = 0x7fffffff;
        for (int j = 1 + lane; j < count; j += 32) {
            int idx = pos + j;
            if (idx >= count) idx -= count;
            short2 pt = src[idx];
            double dx = pt.x - start_pt.x, dy = pt.y - start_pt.y;
            double dist = dx * dx + dy * dy;
            if (dist > best) { best = dist; bj = j; }
        }
        warp_argmax_first(best, bj);
        if (best > 0) right.x = bj;
        le_eps = best <= eps;
    }
    if (!le_eps) {
        right.y = slice.x = pos % count;
        slice.y = right.x = (right.x + slice.x) % count;
        if (lane == 0) { stack[0] = right; stack[1] = slice; }
        top = 2;
    } else {
        if (lane == 0) dst[0] = start_pt;
        new_count = 1;
    }
    __syncwarp();
    while (top > 0) {
        slice = stack[--top];
        end_pt = src[slice.y];
        start_pt = src[slice.x];
        int first = slice.x + 1;
        if (first >= count) first = 0;
        if (first != slice.y) {
            int m = slice.y - first;           // elements first .. slice.y-1, circular
            if (m < 0) m += count;
            double dx = end_pt.x - start_pt.x, dy = end_pt.y - start_pt.y;
            double L = dx * dx + dy * dy, best = 0;
            int bt = 0x7fffffff;
            for (int t = lane; t < m; t += 32) {
                int idx = first + t;
                if (idx >= count) idx -= count;
                short2 pt = src[idx];
                double px = pt.x - start_pt.x, py = pt.y - start_pt.y;
                double tt = px * dx + py * dy, dist;
                if (tt <= 0) dist = (px * px + py * py) * L;
                else if (tt >= L) { double qx = pt.x - end_pt.x, qy = pt.y - end_pt.y; dist = (qx * qx + qy * qy) * L; }
                else { double cr = py * dx - px * dy; dist = cr * cr; }
                if (dist > best) { best = dist; bt = t; }
            }
            warp_argmax_first(best, bt);
            if (best > 0) { int idx = first + bt; right.x = idx >= count ? idx - count : idx; }
            le_eps = best <= eps * L;
        } else {
            le_eps = true;
        }
        __syncwarp();  // everyone has read stack[top] before it may be overwritten
        if (le_eps) {
            if (lane == 0) dst[new_count] = start_pt;
            new_count++;
        } else {
            right.y = slice.y;
            slice.y = right.x;
            if (lane == 0) { stack[top] = right; stack[top + 1] = slice; }
            top += 2;
        }
        __syncwarp();
    }
    if (lane == 0) new_count = approx_cleanup(dst, new_count, eps);
    return __shfl_sync(0xffffffffu, new_count, 0);
}

struct QMeta {  // per frame, zeroed before every batch
    int ncand, nstart, flags, nb, nlong, ncut, fallback, ncc;
    int rewalk, pad_[3];   // rewalk: a piece has more than TMPV vertices (k_qemit_copy -> k_qemit)
};
constexpr int LONG_NP = 96;   // contours with more vertices are handled by a whole warp
constexpr int TMPV = 64;      // vertices of a piece that the first walk keeps; a piece with more is walked again by k_qemit
constexpr int MAXLONG = MANTIS_MAX_PTS / LONG_NP + 1;

struct QuadArgs {
    const uint32_t* morph_bits;
    int W, H, WW, n_frames;
    QMeta* meta;          // [n]
    uint32_t* cand;       // [n][MAXCAND]
    uint32_t* ukey;       // [n][MAXB] unsorted starts: key, vertex count, walk length
    int* unp;
    int* usteps;
    uint32_t* key;        // [n][MAXB] sorted by discovery order
    int* np;
    int* off;
    short* bylen;         // [n][MAXB] border indices, longest walk first
    short* longl;         // [n][MAXLONG] borders with more than LONG_NP vertices
    short* an;            // [n][MAXB] approxPolyDP vertex count
    short2* quad;         // [n][MAXB][4]
    short2* pts;          // [n][MANTIS_MAX_PTS]
    short2* dst;          // [n][MANTIS_MAX_PTS]
    int2* stack;          // [n][MANTIS_MAX_PTS + 2 * MAXB]
    uint16_t* owner;      // [n][W * H]
    uint32_t* cut;        // [n][MAXCUT] cut visits as listed by k_qcut (pixel, side, candidate)
    uint32_t* cut_key;    // [n][MAXCUT] visit keys (pixel << 3 | back direction) + KEY_ flags, from k_qseg
    uint32_t* cut_end;    // [n][MAXCUT] state at which the piece starting here ends
    int* cut_np;          // [n][MAXCUT] CHAIN_APPROX_SIMPLE vertices of the piece
    uint32_t* cut_par;    // [n][MAXCUT] hole candidates: a cut visit of the border that owns the left edge of the run left of the hole
    short* hole_owner;    // [n][MAXB] that border, for the hole borders
    uint32_t* cut_dst;    // [n][MAXCUT] (offset of the piece's vertices in its frame << 10) | border
    short2* cut_tmp;      // [n][MAXCUT][TMPV] the first TMPV vertices of every piece, kept by k_qseg for k_qemit
    int use_cuts;
    int cut_mask;             // cut rows are y & cut_mask == 0 (CUT_ROWS - 1 up to VGA, wider spacing for larger frames)
    const uint32_t* padded;   // zero-framed copies of the planes in global memory (PadImg layout) when a plane does not fit in shared memory
    size_t pad_words;         // words per padded plane
    FrameQuads* out;
    double eps, stretch, radius_mult;
};

__device__ __forceinline__ uint32_t bits_at(const uint32_t* p, int WW, int H, int y, int j)
{
    if (y < 0 || y >= H || j < 0 || j >= WW) return 0u;
    return __ldg(p + (size_t)y * WW + j);
}

// length of the run of ones of `t` starting at bit 0 (0..32)
__device__ __forceinline__ int ones_from0(uint32_t t)
{
    int f = __ffs(~t);
    return f ? f - 1 : 32;
}

// ---- 1. candidates -------------------------------------------------------------------------
// Border-start candidates among the pixels of word j of row y (c = that word): push(is_hole, x).
template <class F>
__device__ __forceinline__ void word_candidates(const uint32_t* bits, int W, int H, int WW, int y, int j, uint32_t c, F push)
{
    const uint32_t cp = bits_at(bits, WW, H, y, j - 1);
    const uint32_t left = (c << 1) | (cp >> 31);
    const uint32_t valid = (j == WW - 1 && (W & 31)) ? ((1u << (W & 31)) - 1u) : 0xffffffffu;
    uint32_t outer = c & ~left;                          // first pixels of foreground runs
    uint32_t hole = (y > 0) ? (~c & left & valid) : 0u;  // first pixels of background runs right of foreground
    if (!(outer | hole)) return;

    auto dilated_above = [&](int jj) {  // foreground of row y-1 at columns p-1, p, p+1, for the columns p of word jj
        uint32_t u = bits_at(bits, WW, H, y - 1, jj);
        return u | (u << 1) | (bits_at(bits, WW, H, y - 1, jj - 1) >> 31) | (u >> 1) | (bits_at(bits, WW, H, y - 1, jj + 1) << 31);
    };
    while (outer) {
        const int b = __ffs(outer) - 1;
        outer &= outer - 1;
        // follow the run to the right; reject as soon as foreground shows up above it
        bool ok = true;
        int jj = j, bb = b;
        uint32_t w = c;
        for (;;) {
            int len = ones_from0(w >> bb);                    // run pixels inside this word: bb .. bb+len-1
            uint32_t m = (len >= 32 ? 0xffffffffu : ((1u << len) - 1u)) << bb;
            if (m & dilated_above(jj)) { ok = false; break; }
            if (bb + len < 32 || jj + 1 >= WW) break;         // run ended inside the word / at the last word
            ++jj; bb = 0;
            w = bits_at(bits, WW, H, y, jj);
            if (!(w & 1u)) break;
        }
        if (ok) push(false, j * 32 + b);
    }
    while (hole) {
        const int b = __ffs(hole) - 1;
        hole &= hole - 1;
        // follow the background run; reject on background above it or on reaching the right frame
        bool ok = true;
        int jj = j, bb = b;
        uint32_t w = ~c;
        for (;;) {
            const uint32_t vj = (jj == WW - 1 && (W & 31)) ? ((1u << (W & 31)) - 1u) : 0xffffffffu;
            int len = ones_from0((w & vj) >> bb);
            uint32_t m = (len >= 32 ? 0xffffffffu : ((1u << len) - 1u)) << bb;
            if (m & ~bits_at(bits, WW, H, y - 1, jj)) { ok = false; break; }
            const int xe = jj * 32 + bb + len;                // first column after the run part
            if (xe >= W) { ok = false; break; }               // background run reaches the frame
            if (bb + len < 32) break;
            ++jj; bb = 0;
            w = ~bits_at(bits, WW, H, y, jj);
            if (!(w & 1u)) break;
        }
        if (ok) push(true, j * 32 + b);
    }
}

__global__ void __launch_bounds__(CNT) k_qcand(QuadArgs a)
{
    const int W = a.W, H = a.H, WW = a.WW, frame = blockIdx.y;
    QMeta* meta = a.meta + frame;
    if (a.use_cuts && !meta->fallback) return;
    const uint32_t* bits = a.morph_bits + (size_t)frame * H * WW;
    for (int i = blockIdx.x * CNT + threadIdx.x; i < H * WW; i += gridDim.x * CNT) {
        const int y = i / WW, j = i - y * WW;
        word_candidates(bits, W, H, WW, y, j, bits_at(bits, WW, H, y, j), [&](bool is_hole, int x) {
            int k = atomicAdd(&meta->ncand, 1);
            if (k < MAXCAND) a.cand[(size_t)frame * MAXCAND + k] = (is_hole ? 0x80000000u : 0u) | ((uint32_t)y << 15) | (uint32_t)x;
            else atomicOr(&meta->flags, FRAME_FLAG_BORDER_OVERFLOW);
        });
    }
}

// ---- 2. probe walks ------------------------------------------------------------------------
// Stages the frame's bit plane, zero-framed, in shared memory (PadImg layout).
__device__ __forceinline__ PadImg stage_padded(const uint32_t* g, int H, int WW)
{
    extern __shared__ uint32_t s_bits[];
    const int stride = WW + 2, total = (H + 2) * stride;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int r = i / stride, cidx = i - r * stride;
        s_bits[i] = (r >= 1 && r <= H && cidx >= 1 && cidx <= WW) ? __ldg(g + (size_t)(r - 1) * WW + (cidx - 1)) : 0u;
    }
    __syncthreads();
    PadImg im;
    im.base = s_bits;
    im.stride = stride;
    return im;
}

// The same layout in global memory, for frames whose plane does not fit in shared memory (1920x1080): the
// walkers then read it through L1 / L2 (268 KB per frame).
__device__ __forceinline__ PadImg global_padded(const QuadArgs& a, int frame)
{
    PadImg im;
    im.base = a.padded + (size_t)frame * a.pad_words;
    im.stride = a.WW + 2;
    return im;
}

__global__ void __launch_bounds__(CNT) k_qpad(QuadArgs a, uint32_t* dst)
{
    const int frame = blockIdx.y, H = a.H, WW = a.WW;
    const uint32_t* g = a.morph_bits + (size_t)frame * H * WW;
    uint32_t* d = dst + (size_t)frame * a.pad_words;
    const int stride = WW + 2, total = (H + 2) * stride;
    for (int i = blockIdx.x * CNT + threadIdx.x; i < total; i += gridDim.x * CNT) {
        const int r = i / stride, cidx = i - r * stride;
        d[i] = (r >= 1 && r <= H && cidx >= 1 && cidx <= WW) ? __ldg(g + (size_t)(r - 1) * WW + (cidx - 1)) : 0u;
    }
}

template <class Img>
__device__ __forceinline__ void probe_frame(const QuadArgs& a, const Img& im, int frame)
{
    QMeta* meta = a.meta + frame;
    const int ncand = min(meta->ncand, MAXCAND);
    const int W = a.W, H = a.H;
    for (int i = threadIdx.x; i < ncand; i += CNT) {
        const uint32_t cand = a.cand[(size_t)frame * MAXCAND + i];
        const bool is_hole = cand >> 31;
        const int y = (cand >> 15) & 0xffff, x = cand & 0x7fff;
        int np, steps;
        bool ok;
        if (!is_hole) {
            OuterProbe v{x, y, 0, 0};
            ok = follow_border(im, x, y, false, v);
            np = v.npts; steps = v.steps;
        } else {
            HoleProbe v{x, y, W, H, 0, 0};
            ok = follow_border(im, x - 1, y, true, v);
            np = v.npts; steps = v.steps;
        }
        if (!ok) continue;
        int k = atomicAdd(&meta->nstart, 1);
        if (k >= MAXB) { atomicOr(&meta->flags, FRAME_FLAG_BORDER_OVERFLOW); continue; }
        const size_t o = (size_t)frame * MAXB + k;
        a.ukey[o] = ((uint32_t)(y * (W + 1) + x) << 1) | (is_hole ? 1u : 0u);
        a.unp[o] = np;
        a.usteps[o] = steps;
    }
}

__global__ void __launch_bounds__(CNT) k_qprobe(QuadArgs a, int use_smem)
{
    const int frame = blockIdx.x;
    if (a.use_cuts && !a.meta[frame].fallback) return;
    const uint32_t* g = a.morph_bits + (size_t)frame * a.H * a.WW;
    if (use_smem == 1) {
        probe_frame(a, stage_padded(g, a.H, a.WW), frame);
    } else if (use_smem == 2) {
        probe_frame(a, global_padded(a, frame), frame);
    } else {
        BitImg im;
        im.bits = g; im.W = a.W; im.H = a.H; im.WW = a.WW;
        probe_frame(a, im, frame);
    }
}

// ---- 3. discovery order, point offsets, by-length order ---------------------------------------
struct RankSmem {
    uint32_t ukey[MAXB];
    int ust[MAXB], np[MAXB], st[MAXB];
};
// Ranks the frame's starts a.ukey/unp/usteps[0..nb) (global).  rank_of (shared, may be null)
// receives the discovery index of every unsorted start.
__device__ void rank_frame(const QuadArgs& a, int frame, RankSmem& r, int nb, short* rank_of)
{
    const int tid = threadIdx.x, nt = blockDim.x;
    QMeta* meta = a.meta + frame;
    const size_t base = (size_t)frame * MAXB;
    for (int i = tid; i < nb; i += nt) { r.ukey[i] = a.ukey[base + i]; r.ust[i] = a.usteps[base + i]; }
    __syncthreads();
    for (int i = tid; i < nb; i += nt) {
        const uint32_t k = r.ukey[i];
        int rk = 0;
        for (int j = 0; j < nb; ++j) rk += r.ukey[j] < k;
        a.key[base + rk] = k;
        r.np[rk] = a.unp[base + i];
        r.st[rk] = r.ust[i];
        if (rank_of) rank_of[i] = (short)rk;
    }
    __syncthreads();
    if (tid == 0) {
        int acc = 0;
        for (int i = 0; i < nb; ++i) { a.off[base + i] = acc; acc += r.np[i]; }
        if (acc > MANTIS_MAX_PTS) meta->flags |= FRAME_FLAG_POINT_OVERFLOW;
        meta->nb = nb;
    }
    for (int i = tid; i < nb; i += nt) {
        a.np[base + i] = r.np[i];
        const int st = r.st[i];
        int rk = 0;  // position by descending walk length (ties by index)
        for (int j = 0; j < nb; ++j) rk += (r.st[j] > st) || (r.st[j] == st && j < i);
        a.bylen[base + rk] = (short)i;
        if (r.np[i] > LONG_NP) {
            int k = atomicAdd(&meta->nlong, 1);
            if (k < MAXLONG) a.longl[(size_t)frame * MAXLONG + k] = (short)i;
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(CNT) k_qrank(QuadArgs a)
{
    __shared__ RankSmem r;
    const int frame = blockIdx.x;
    if (a.use_cuts && !a.meta[frame].fallback) return;
    rank_frame(a, frame, r, min(a.meta[frame].nstart, MAXB), nullptr);
}

// ---- 4. extraction + approxPolyDP, one thread per border --------------------------------------
constexpr int TNT = 256;
template <class Img>
__device__ __forceinline__ void trace_frame(const QuadArgs& a, const Img& im, int frame)
{
    const QMeta* meta = a.meta + frame;
    if (meta->flags) return;
    const int nb = meta->nb, W = a.W, H = a.H;
    const size_t base = (size_t)frame * MAXB;
    for (int slot = threadIdx.x; slot < nb; slot += TNT) {
        const int b = a.bylen[base + slot];
        const uint32_t k = a.key[base + b];
        const bool is_hole = k & 1u;
        const int p = (int)(k >> 1), y = p / (W + 1), x = p - y * (W + 1);
        const int off = a.off[base + b], np = a.np[base + b];
        short2* pts = a.pts + (size_t)frame * MANTIS_MAX_PTS + off;
        short2* dst = a.dst + (size_t)frame * MANTIS_MAX_PTS + off;
        int2* stack = a.stack + (size_t)frame * (MANTIS_MAX_PTS + 2 * MAXB) + off + 2 * b;
        Emit v{pts, a.owner + (size_t)frame * W * H, W, 0, (uint16_t)b};
        follow_border(im, is_hole ? x - 1 : x, y, is_hole, v);
        if (np > LONG_NP) continue;  // approximated by a whole warp in k_qapprox_long
        const int n4 = approx_poly_closed(pts, np, a.eps, dst, stack);
        a.an[base + b] = (short)min(n4, 32767);
        if (n4 == 4)
            for (int q = 0; q < 4; ++q) a.quad[(base + b) * 4 + q] = dst[q];
    }
}

__global__ void __launch_bounds__(TNT) k_qtrace(QuadArgs a, int use_smem)
{
    const int frame = blockIdx.x;
    if (a.use_cuts && !a.meta[frame].fallback) return;
    const uint32_t* g = a.morph_bits + (size_t)frame * a.H * a.WW;
    if (use_smem == 1) {
        if (a.meta[frame].flags || a.meta[frame].nb == 0) return;
        trace_frame(a, stage_padded(g, a.H, a.WW), frame);
    } else if (use_smem == 2) {
        if (a.meta[frame].flags || a.meta[frame].nb == 0) return;
        trace_frame(a, global_padded(a, frame), frame);
    } else {
        BitImg im;
        im.bits = g; im.W = a.W; im.H = a.H; im.WW = a.WW;
        trace_frame(a, im, frame);
    }
}

// ---- segment-parallel tracing -----------------------------------------------------------------
// A border is a cycle of VISITS (pixel, direction back to the previous pixel); a visit examines one
// maximal run of background neighbours, and every background 4-neighbour of a foreground pixel is
// examined by exactly one visit of exactly one border.  A visit can therefore be recognised from its
// pixel's 3x3 neighbourhood alone, and borders can be cut at visits that are listed without walking:
//   candidates  the visit examining West of an outer-border candidate pixel, and the visit
//               examining East of the pixel left of a hole candidate (the run tests of k_qcand);
//               a border's start is the raster-first candidate visit of its cycle, and that
//               candidate's kind says whether the cycle is an outer or a hole border,
//   rows        every visit examining West or East on the rows y % CUT_ROWS == 0 (bounds the pieces).
//   k_qcut   lists the cut visits of a frame, bit-parallel.
//   k_qseg   walks every piece to the next cut (lanes take pieces from a shared counter), counting its
//            CHAIN_APPROX_SIMPLE vertices and keeping the first TMPV of them.
//   k_qlink  per frame: links the pieces into cycles (hash of visit keys), finds every cycle's
//            start, ranks the starts in discovery order, gives every piece its place in the vertex array.
//   k_qemit_copy  puts the kept vertices at their place; k_qemit walks the few pieces with more than TMPV again.
// approxPolyDP then runs per border (k_qapprox_long).  A frame that exceeds MAXCUT / MAXCC, or whose
// links do not close, is handed to the sequential-walk kernels above (meta->fallback).
constexpr int CUT_ROWS = 16;
constexpr int MAXCUT = 2048;   // cut visits per frame
constexpr int MAXCC = 512;     // candidates among them
constexpr int TABN = 2048;     // candidate table (open addressing)
constexpr int HASHN = 4096;
constexpr uint32_t KEY_LTOP = 1u << 30, KEY_RCAND = 1u << 31, KEY_STATE = 0x00ffffffu, KEY_DEAD = 0x0f000000u;
constexpr uint32_t EMPTY = 0xffffffffu;

// candidate visits of the frame, keyed (pixel << 1) | (1 if it is the East side)
struct CutSet {
    uint32_t* tab;  // [TABN], shared
    int W, mask;
    static __device__ __forceinline__ uint32_t slot(uint32_t k) { return (k * 2654435761u) >> 21; }  // TABN = 2^11
    __device__ __forceinline__ void insert(uint32_t k)
    {
        for (uint32_t h = slot(k);; h = (h + 1) & (TABN - 1)) {
            const uint32_t old = atomicCAS(&tab[h], EMPTY, k);
            if (old == EMPTY || old == k) return;
        }
    }
    __device__ __forceinline__ bool has(uint32_t k) const
    {
        for (uint32_t h = slot(k);; h = (h + 1) & (TABN - 1)) {
            const uint32_t t = tab[h];
            if (t == k) return true;
            if (t == EMPTY) return false;
        }
    }
    __device__ __forceinline__ bool cand_west(int x, int y, uint32_t nb) const { return (nb & 0x0eu) == 0u && has((uint32_t)(y * W + x) << 1); }
    __device__ __forceinline__ bool cand_east(int x, int y, uint32_t nb) const { return (nb & 0x02u) != 0u && has(((uint32_t)(y * W + x) << 1) | 1u); }
    __device__ __forceinline__ bool cut_west(int x, int y, uint32_t nb) const { return (y & mask) == 0 || cand_west(x, y, nb); }
    __device__ __forceinline__ bool cut_east(int x, int y, uint32_t nb) const { return (y & mask) == 0 || cand_east(x, y, nb); }
    // is the visit at (x, y) that examined the background directions z a cut?
    __device__ __forceinline__ bool is_cut(int x, int y, uint32_t nb, uint32_t z) const
    {
        return ((z & 0x10u) && cut_west(x, y, nb)) || ((z & 0x01u) && cut_east(x, y, nb));
    }
};

// raw entries: (pixel << 2) | (2 if East side) | (1 if candidate)
__global__ void __launch_bounds__(CNT) k_qcut(QuadArgs a)
{
    const int W = a.W, H = a.H, WW = a.WW, frame = blockIdx.y;
    const uint32_t* bits = a.morph_bits + (size_t)frame * H * WW;
    QMeta* meta = a.meta + frame;
    uint32_t* dst = a.cut + (size_t)frame * MAXCUT;
    for (int i = blockIdx.x * CNT + threadIdx.x; i < H * WW; i += gridDim.x * CNT) {
        const uint32_t c = __ldg(bits + i);
        const int y = i / WW, j = i - y * WW;
        word_candidates(bits, W, H, WW, y, j, c, [&](bool is_hole, int x) {
            const int k = atomicAdd(&meta->ncut, 1);
            if (atomicAdd(&meta->ncc, 1) >= MAXCC || k >= MAXCUT) { meta->fallback = 1; return; }
            dst[k] = is_hole ? ((((uint32_t)(y * W + x - 1)) << 2) | 3u) : ((((uint32_t)(y * W + x)) << 2) | 1u);
        });
        if (!c || (y & a.cut_mask)) continue;
        const uint32_t cp = bits_at(bits, WW, H, y, j - 1), cn = bits_at(bits, WW, H, y, j + 1);
        uint32_t west = c & ~((c << 1) | (cp >> 31));   // foreground with background to the West
        uint32_t east = c & ~((c >> 1) | (cn << 31));   // ... to the East
        const int n = __popc(west) + __popc(east);
        int k = atomicAdd(&meta->ncut, n);
        if (k + n > MAXCUT) { meta->fallback = 1; continue; }
        const uint32_t pos0 = (uint32_t)(y * W + j * 32);
        while (west) { const int b = __ffs(west) - 1; west &= west - 1; dst[k++] = (pos0 + b) << 2; }
        while (east) { const int b = __ffs(east) - 1; east &= east - 1; dst[k++] = ((pos0 + b) << 2) | 2u; }
    }
}

struct SegKeep {   // counts the vertices of a piece and keeps the first TMPV of them (out may be null: count only)
    short2* out;
    int np;
    __device__ __forceinline__ void visit(int, int, uint32_t) {}
    __device__ __forceinline__ void point(int x, int y)
    {
        if (out && np < TMPV) out[np] = make_short2((short)x, (short)y);
        ++np;
    }
};
struct SegEmit {
    short2* out;
    __device__ __forceinline__ void visit(int, int, uint32_t) {}
    __device__ __forceinline__ void point(int x, int y) { *out++ = make_short2((short)x, (short)y); }
};

// Walks from the visit (x, y, s) -- s = direction back to the previous pixel, nb = neighbour mask of
// (x, y), not empty -- to the next cut visit and returns that visit's key.  The steps are those of
// follow_border; the first visit is processed without asking whether it is a cut.
template <class Img, class V>
__device__ __forceinline__ uint32_t walk_piece(const Img& im, const CutSet& cs, int x, int y, int s, uint32_t nb, V& vis)
{
    int prev_s = s ^ 4;
    bool first = true;
    for (int guard = 0; guard < (1 << 22); ++guard) {
        const int r0 = (s + 1) & 7;
        const uint32_t rot = ((nb | (nb << 8)) >> r0) & 0xffu;
        const int k = __ffs(rot) - 1;
        uint32_t z = ((1u << k) - 1u) << r0;
        z = (z | (z >> 8)) & 0xffu;
        if (!first && cs.is_cut(x, y, nb, z)) return ((uint32_t)(y * cs.W + x) << 3) | (uint32_t)s;
        first = false;
        s = (r0 + k) & 7;
        vis.visit(x, y, z);
        if (s != prev_s) {
            vis.point(x, y);
            prev_s = s;
        }
        x += dir_dx(s);
        y += dir_dy(s);
        s = (s + 4) & 7;
        nb = im.nbr(x, y);
    }
    return EMPTY;  // no cut met: k_qlink finds no such visit and hands the frame to the sequential route
}

// direction of the first foreground neighbour clockwise from `side` (nb not empty, `side` background)
__device__ __forceinline__ int back_dir(uint32_t nb, int side)
{
    const uint32_t rot = ((nb | (nb << 8)) >> side) & 0xffu;  // bit j: direction side + j
    return (side + 31 - __clz((int)rot)) & 7;
}

// One step of walk_piece() with the walk's state in the caller's registers (the lanes of a warp are each inside a
// different piece).  Returns true when (x, y, s) is the cut visit that ends the piece; else moves on one border pixel.
template <class Img>
__device__ __forceinline__ bool walk_step(const Img& im, const CutSet& cs, int& x, int& y, int& s, int& prev_s, uint32_t& nb,
                                          bool& first, int& np, short2* tmp)
{
    const int r0 = (s + 1) & 7;
    const uint32_t rot = ((nb | (nb << 8)) >> r0) & 0xffu;
    const int k = __ffs(rot) - 1;
    uint32_t z = ((1u << k) - 1u) << r0;
    z = (z | (z >> 8)) & 0xffu;
    if (!first && cs.is_cut(x, y, nb, z)) return true;
    first = false;
    s = (r0 + k) & 7;
    if (s != prev_s) {
        if (tmp && np < TMPV) tmp[np] = make_short2((short)x, (short)y);
        ++np;
        prev_s = s;
    }
    x += dir_dx(s);
    y += dir_dy(s);
    s = (s + 4) & 7;
    nb = im.nbr(x, y);
    return false;
}

// One CTA per frame.  Pieces are 1 .. 400 steps long (19 on average, measured), so with a fixed piece per lane a
// warp waits for the longest of its 32: a third of the lanes work.  Here a lane takes the next cut of the frame
// from a shared counter whenever it has finished its piece (checked every STEPS steps; same scheme as pose.cu's
// k_pose_run).  The hole candidates' walks to a cut of their parent border run the same way in a second pass.
template <bool GPAD>
__global__ void __launch_bounds__(CNT) k_qseg(QuadArgs a)
{
    __shared__ uint32_t s_tab[TABN];
    __shared__ int s_next, s_npar, s_next_par;
    __shared__ unsigned short s_par[MAXCC];   // slots whose hole needs a parent walk
    const int frame = blockIdx.x, W = a.W, lane = threadIdx.x & 31;
    QMeta* meta = a.meta + frame;
    if (meta->fallback) return;
    const int ncut = meta->ncut;
    const size_t base = (size_t)frame * MAXCUT;
    CutSet cs{s_tab, W, a.cut_mask};
    for (int i = threadIdx.x; i < TABN; i += CNT) s_tab[i] = EMPTY;
    if (threadIdx.x == 0) { s_next = 0; s_npar = 0; s_next_par = 0; }
    __syncthreads();
    for (int i = threadIdx.x; i < ncut; i += CNT) {
        const uint32_t raw = a.cut[base + i];
        if (raw & 1u) cs.insert(raw >> 1);
    }
    PadImg im;
    if (GPAD) { im = global_padded(a, frame); __syncthreads(); }
    else im = stage_padded(a.morph_bits + (size_t)frame * a.H * a.WW, a.H, a.WW);  // (synchronises)

    constexpr unsigned FULL = 0xffffffffu;
    constexpr int STEPS = 8, GUARD = 1 << 22;   // steps after which a walk is given up (walk_piece's bound)
    // ---- pass 1: the pieces ----
    {
        int slot = 0, x = 0, y = 0, s = 0, prev_s = 0, np = 0, steps = 0;
        uint32_t nb = 0, key = 0;
        bool first = true, active = false, more = true, finished = false;
        uint32_t end_key = 0;
        short2* tmp = nullptr;
        for (;;) {
            if (finished) {   // results of the piece that ended during the last STEPS steps (kept out of the step loop)
                a.cut_key[base + slot] = key;
                a.cut_end[base + slot] = end_key;
                a.cut_np[base + slot] = np;
                if (np > 65535) meta->fallback = 1;
                if (key & KEY_RCAND) {
                    const int k = atomicAdd(&s_npar, 1);
                    if (k < MAXCC) s_par[k] = (unsigned short)slot; else meta->fallback = 1;
                }
                finished = false;
            }
            const unsigned want = __ballot_sync(FULL, !active && more);
            if (want) {
                const int leader = __ffs(want) - 1;
                int first_slot = 0;
                if (lane == leader) first_slot = atomicAdd(&s_next, __popc(want));
                first_slot = __shfl_sync(FULL, first_slot, leader);
                if (!active && more) {
                    slot = first_slot + __popc(want & ((1u << lane) - 1u));
                    if (slot >= ncut) {
                        more = false;
                    } else {
                        const uint32_t raw = a.cut[base + slot];
                        const bool east = raw & 2u;
                        const int side = east ? 0 : 4, pos = (int)(raw >> 2);
                        y = pos / W; x = pos - y * W;
                        nb = im.nbr(x, y);
                        key = KEY_DEAD;
                        uint32_t end = 0;
                        np = 0;
                        if (!(raw & 1u) && cs.has(raw >> 1)) {
                            // a row cut that is also listed as a candidate: the candidate entry stands for it
                        } else if (nb == 0u) {  // isolated pixel: one visit, one vertex; stands under its West side
                            if (!east) {
                                key = ((uint32_t)pos << 3) | KEY_LTOP; end = (uint32_t)pos << 3; np = 1;
                                a.cut_tmp[(base + slot) * TMPV] = make_short2((short)x, (short)y);
                            }
                        } else {
                            s = back_dir(nb, side);
                            const int r0 = (s + 1) & 7;
                            const uint32_t rot = ((nb | (nb << 8)) >> r0) & 0xffu;
                            uint32_t z = ((1u << (__ffs(rot) - 1)) - 1u) << r0;
                            z = (z | (z >> 8)) & 0xffu;
                            if (!(east && (z & 0x10u) && cs.cut_west(x, y, nb))) {  // else the same visit stands under its West side
                                key = ((uint32_t)pos << 3) | (uint32_t)s;
                                if ((z & 0x10u) && cs.cand_west(x, y, nb)) key |= KEY_LTOP;
                                if ((z & 0x01u) && cs.cand_east(x, y, nb)) key |= KEY_RCAND;
                                prev_s = s ^ 4; first = true; steps = 0;
                                tmp = a.cut_tmp + (base + slot) * TMPV;
                                active = true;
                            }
                        }
                        if (!active) {
                            a.cut_key[base + slot] = key;
                            a.cut_end[base + slot] = end;
                            a.cut_np[base + slot] = np;
                        }
                    }
                }
            }
            if (!__any_sync(FULL, active || more)) break;
#pragma unroll 1
            for (int i = 0; i < STEPS; ++i) {
                if (!active) continue;
                const bool done = walk_step(im, cs, x, y, s, prev_s, nb, first, np, tmp);
                if (!done && ++steps < GUARD) continue;
                // EMPTY: no cut met -- k_qlink finds no such visit and hands the frame to the sequential route
                end_key = done ? (((uint32_t)(y * W + x) << 3) | (uint32_t)s) : EMPTY;
                active = false;
                finished = true;
            }
        }
    }
    __syncthreads();
    // ---- pass 2: RETR_CCOMP parent of the hole that would start at a KEY_RCAND visit: the border passing the left
    // edge of the foreground run that ends at this pixel.  Any cut visit of that border identifies it: the visit
    // examining that left edge, or the first cut after it. ----
    {
        const int npar = min(s_npar, MAXCC);
        int slot = 0, x = 0, y = 0, s = 0, prev_s = 0, np = 0, steps = 0;
        uint32_t nb = 0;
        bool first = true, active = false, more = true, finished = false;
        uint32_t end_key = 0;
        for (;;) {
            if (finished) { a.cut_par[base + slot] = end_key; finished = false; }
            const unsigned want = __ballot_sync(FULL, !active && more);
            if (want) {
                const int leader = __ffs(want) - 1;
                int first_item = 0;
                if (lane == leader) first_item = atomicAdd(&s_next_par, __popc(want));
                first_item = __shfl_sync(FULL, first_item, leader);
                if (!active && more) {
                    const int item = first_item + __popc(want & ((1u << lane) - 1u));
                    if (item >= npar) {
                        more = false;
                    } else {
                        slot = s_par[item];
                        const int pos = (int)(a.cut[base + slot] >> 2);
                        y = pos / W;
                        x = pos - y * W;
                        while (im.at(x - 1, y)) --x;
                        nb = im.nbr(x, y);  // not empty: the run's right end has a neighbour above
                        s = back_dir(nb, 4);
                        const int r0 = (s + 1) & 7;
                        const uint32_t rot = ((nb | (nb << 8)) >> r0) & 0xffu;
                        uint32_t z = ((1u << (__ffs(rot) - 1)) - 1u) << r0;
                        z = (z | (z >> 8)) & 0xffu;
                        if (cs.is_cut(x, y, nb, z)) {
                            a.cut_par[base + slot] = ((uint32_t)(y * W + x) << 3) | (uint32_t)s;
                        } else {
                            prev_s = s ^ 4; first = true; steps = 0; np = 0;
                            active = true;
                        }
                    }
                }
            }
            if (!__any_sync(FULL, active || more)) break;
#pragma unroll 1
            for (int i = 0; i < STEPS; ++i) {
                if (!active) continue;
                const bool done = walk_step(im, cs, x, y, s, prev_s, nb, first, np, nullptr);
                if (!done && ++steps < GUARD) continue;
                end_key = done ? (((uint32_t)(y * W + x) << 3) | (uint32_t)s) : EMPTY;
                active = false;
                finished = true;
            }
        }
    }
}

constexpr int LNT = 256;
__global__ void __launch_bounds__(LNT) k_qlink(QuadArgs a)
{
    __shared__ uint32_t s_key[MAXCUT];
    __shared__ uint16_t s_succ[MAXCUT], s_np[MAXCUT];
    __shared__ short s_start[MAXB], s_rank[MAXB], s_par[MAXB];
    __shared__ union U {
        struct { uint32_t key[HASHN]; uint16_t val[HASHN]; } h;
        RankSmem r;
        __device__ U() {}
    } u;
    __shared__ int s_bad, s_nstart;

    const int frame = blockIdx.x, tid = threadIdx.x, W = a.W;
    QMeta* meta = a.meta + frame;
    if (meta->fallback) return;
    const int ncut = meta->ncut;
    const size_t base = (size_t)frame * MAXCUT, bbase = (size_t)frame * MAXB;
    for (int i = tid; i < HASHN; i += LNT) u.h.key[i] = EMPTY;
    for (int i = tid; i < ncut; i += LNT) {
        s_key[i] = a.cut_key[base + i];
        s_np[i] = (uint16_t)a.cut_np[base + i];
        a.cut_dst[base + i] = EMPTY;
    }
    if (tid == 0) { s_bad = 0; s_nstart = 0; }
    __syncthreads();
    auto slot_of = [](uint32_t st) { return (st * 2654435761u) >> 20; };  // HASHN = 2^12
    for (int i = tid; i < ncut; i += LNT) {
        const uint32_t k = s_key[i];
        if (k & KEY_DEAD) continue;
        const uint32_t st = k & KEY_STATE;
        for (uint32_t h = slot_of(st);; h = (h + 1) & (HASHN - 1)) {
            const uint32_t old = atomicCAS(&u.h.key[h], EMPTY, st);
            if (old == EMPTY) { u.h.val[h] = (uint16_t)i; break; }
            if (old == st) { s_bad = 1; break; }  // one visit listed twice: not expected
        }
    }
    __syncthreads();
    for (int i = tid; i < ncut; i += LNT) {
        if (s_key[i] & KEY_DEAD) continue;
        const uint32_t st = a.cut_end[base + i];
        uint32_t h = slot_of(st);
        for (;; h = (h + 1) & (HASHN - 1)) {
            const uint32_t k = u.h.key[h];
            if (k == st || k == EMPTY) break;
        }
        if (u.h.key[h] == st) s_succ[i] = u.h.val[h];
        else s_bad = 1;  // the piece ended at a visit nobody listed
    }
    __syncthreads();
    if (s_bad) { if (tid == 0) meta->fallback = 1; return; }
    // every cycle's start: its raster-first candidate visit
    for (int i = tid; i < ncut; i += LNT) {
        const uint32_t k = s_key[i];
        if ((k & KEY_DEAD) || !(k & (KEY_LTOP | KEY_RCAND))) continue;
        const uint32_t mine = k & KEY_STATE;
        int total = s_np[i], hops = 0;
        bool first = true;
        for (int j = s_succ[i]; j != i; j = s_succ[j]) {
            const uint32_t kj = s_key[j];
            if ((kj & (KEY_LTOP | KEY_RCAND)) && (kj & KEY_STATE) < mine) { first = false; break; }
            total += s_np[j];
            if (++hops > ncut) { s_bad = 1; first = false; break; }
        }
        if (!first) continue;
        const int n = atomicAdd(&s_nstart, 1);
        if (n >= MAXB) continue;
        const int pos = (int)(mine >> 3), y = pos / W, x = pos - y * W;
        const bool hole = k & KEY_RCAND;
        s_start[n] = (short)i;
        if (hole) {  // a piece of the border that owns the hole (its parent in RETR_CCOMP, or a hole beside it)
            const uint32_t st = a.cut_par[base + i];
            uint32_t h = slot_of(st);
            for (;; h = (h + 1) & (HASHN - 1)) {
                const uint32_t kk = u.h.key[h];
                if (kk == st || kk == EMPTY) break;
            }
            if (u.h.key[h] == st) s_par[n] = (short)u.h.val[h];
            else s_bad = 1;
        }
        a.ukey[bbase + n] = ((uint32_t)(y * (W + 1) + x + (hole ? 1 : 0)) << 1) | (hole ? 1u : 0u);
        a.unp[bbase + n] = total;
        a.usteps[bbase + n] = total;
    }
    __syncthreads();
    if (s_bad) { if (tid == 0) meta->fallback = 1; return; }
    const int nstart = s_nstart;
    if (nstart > MAXB) {
        if (tid == 0) { meta->nstart = nstart; meta->flags |= FRAME_FLAG_BORDER_OVERFLOW; }
        return;
    }
    if (tid == 0) meta->nstart = nstart;
    rank_frame(a, frame, u.r, nstart, s_rank);
    if (meta->flags) return;
    for (int n = tid; n < nstart; n += LNT) {
        const int b = s_rank[n], i0 = s_start[n];
        int run = a.off[bbase + b], j = i0;
        do {
            a.cut_dst[base + j] = ((uint32_t)run << 10) | (uint32_t)b;
            run += s_np[j];
            j = s_succ[j];
        } while (j != i0);
    }
    __syncthreads();
    for (int n = tid; n < nstart; n += LNT)
        if (s_key[s_start[n]] & KEY_RCAND) a.hole_owner[bbase + s_rank[n]] = (short)(a.cut_dst[base + s_par[n]] & 1023u);
}

// Vertices to their place in the frame's vertex array.  k_qemit_copy: what k_qseg kept for pieces of up to TMPV
// vertices (all but a handful): no shared memory, so eight CTAs per SM wait on the loads instead of four, and the
// first eight vertices of a piece travel as two 16-byte loads issued together (the kernel is nothing but memory
// latency).  A frame with a longer piece is flagged and k_qemit walks those pieces again.
__global__ void __launch_bounds__(CNT) k_qemit_copy(QuadArgs a)
{
    const int frame = blockIdx.x;
    QMeta* meta = a.meta + frame;
    if (meta->fallback || meta->flags || meta->nb == 0) return;
    const int ncut = meta->ncut;
    const size_t base = (size_t)frame * MAXCUT;
    for (int slot = threadIdx.x; slot < ncut; slot += CNT) {
        const uint32_t d = a.cut_dst[base + slot];
        const int np = a.cut_np[base + slot];
        if (d == EMPTY) continue;
        if (np > TMPV) { meta->rewalk = 1; continue; }
        const uint4* src = reinterpret_cast<const uint4*>(a.cut_tmp + (base + slot) * TMPV);   // 256-byte aligned
        uint32_t* dst = reinterpret_cast<uint32_t*>(a.pts + (size_t)frame * MANTIS_MAX_PTS + (d >> 10));
        uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
        if (np > 0) v0 = src[0];
        if (np > 4) v1 = src[1];
        if (np > 0) dst[0] = v0.x;
        if (np > 1) dst[1] = v0.y;
        if (np > 2) dst[2] = v0.z;
        if (np > 3) dst[3] = v0.w;
        if (np > 4) dst[4] = v1.x;
        if (np > 5) dst[5] = v1.y;
        if (np > 6) dst[6] = v1.z;
        if (np > 7) dst[7] = v1.w;
        const uint32_t* s1 = reinterpret_cast<const uint32_t*>(src);
        for (int k = 8; k < np; ++k) dst[k] = s1[k];
    }
}

template <bool GPAD>
__global__ void __launch_bounds__(CNT) k_qemit(QuadArgs a)
{
    __shared__ uint32_t s_tab[TABN];
    const int frame = blockIdx.x, W = a.W;
    const QMeta* meta = a.meta + frame;
    if (meta->fallback || meta->flags || meta->nb == 0 || !meta->rewalk) return;
    const int ncut = meta->ncut;
    const size_t base = (size_t)frame * MAXCUT;
    CutSet cs{s_tab, W, a.cut_mask};
    for (int i = threadIdx.x; i < TABN; i += CNT) s_tab[i] = EMPTY;
    __syncthreads();
    for (int i = threadIdx.x; i < ncut; i += CNT) {
        const uint32_t key = a.cut_key[base + i];
        if (key & KEY_DEAD) continue;
        if (key & KEY_LTOP) cs.insert(((key & KEY_STATE) >> 3) << 1);
        if (key & KEY_RCAND) cs.insert((((key & KEY_STATE) >> 3) << 1) | 1u);
    }
    PadImg im;
    if (GPAD) { im = global_padded(a, frame); __syncthreads(); }
    else im = stage_padded(a.morph_bits + (size_t)frame * a.H * a.WW, a.H, a.WW);  // (synchronises)
    for (int slot = threadIdx.x; slot < ncut; slot += CNT) {
        const uint32_t d = a.cut_dst[base + slot];
        if (d == EMPTY || a.cut_np[base + slot] <= TMPV) continue;
        const uint32_t key = a.cut_key[base + slot];
        const int pos = (int)((key & KEY_STATE) >> 3), y = pos / W, x = pos - y * W;
        SegEmit v{a.pts + (size_t)frame * MANTIS_MAX_PTS + (d >> 10)};
        walk_piece(im, cs, x, y, (int)(key & 7u), im.nbr(x, y), v);
    }
}

// long contours: approxPolyDP by a whole warp
constexpr int LWARPS = 4;
__global__ void __launch_bounds__(32 * LWARPS) k_qapprox_long(QuadArgs a)
{
    const int frame = blockIdx.x, lane = threadIdx.x & 31;
    const QMeta* meta = a.meta + frame;
    const int nlong = min(meta->nlong, MAXLONG);
    if (meta->flags) return;
    const size_t base = (size_t)frame * MAXB;
    for (int k = threadIdx.x >> 5; k < nlong; k += LWARPS) {
        const int b = a.longl[(size_t)frame * MAXLONG + k];
        const int off = a.off[base + b], np = a.np[base + b];
        short2* pts = a.pts + (size_t)frame * MANTIS_MAX_PTS + off;
        short2* dst = a.dst + (size_t)frame * MANTIS_MAX_PTS + off;
        int2* stack = a.stack + (size_t)frame * (MANTIS_MAX_PTS + 2 * MAXB) + off + 2 * b;
        const int n4 = approx_poly_closed_warp(pts, np, a.eps, dst, stack);
        if (lane == 0) {
            a.an[base + b] = (short)min(n4, 32767);
            if (n4 == 4)
                for (int q = 0; q < 4; ++q) a.quad[(base + b) * 4 + q] = dst[q];
        }
        __syncwarp();
    }
    if (!a.use_cuts || meta->fallback) return;  // the sequential route approximates short borders while tracing
    const int nb = meta->nb;
    for (int slot = threadIdx.x; slot < nb; slot += 32 * LWARPS) {
        const int b = a.bylen[base + slot];
        const int off = a.off[base + b], np = a.np[base + b];
        if (np > LONG_NP) continue;
        short2* dst = a.dst + (size_t)frame * MANTIS_MAX_PTS + off;
        const int n4 = approx_poly_closed(a.pts + (size_t)frame * MANTIS_MAX_PTS + off, np, a.eps, dst,
                                          a.stack + (size_t)frame * (MANTIS_MAX_PTS + 2 * MAXB) + off + 2 * b);
        a.an[base + b] = (short)min(n4, 32767);
        if (n4 == 4)
            for (int q = 0; q < 4; ++q) a.quad[(base + b) * 4 + q] = dst[q];
    }
}

// ---- 5. hole parents, CCOMP order, quads, de-duplication, output (per frame) --------------------
constexpr int FNT = 128;
__global__ void __launch_bounds__(FNT) k_qfinish(QuadArgs a)
{
    __shared__ uint32_t s_key[MAXB];
    __shared__ short s_parent[MAXB];
    __shared__ short s_pos2b[MAXB];
    __shared__ short s_an[MAXB];
    __shared__ float s_cx[MANTIS_MAX_QUADS], s_cy[MANTIS_MAX_QUADS];
    __shared__ short s_qb[MANTIS_MAX_QUADS];
    __shared__ unsigned char s_neighbor[MANTIS_MAX_QUADS];
    __shared__ int s_nraw, s_qflag;

    const int frame = blockIdx.x, tid = threadIdx.x;
    const int W = a.W, H = a.H;
    const QMeta* meta = a.meta + frame;
    FrameQuads* out = a.out + frame;
    const int flags = meta->flags;
    const int nb = flags ? min(meta->nstart, MAXB) : meta->nb;
    if (flags) {
        if (tid == 0) { out->n_contours = nb; out->flags = flags; out->n_quads_raw = 0; out->n_quads = 0; }
        return;
    }
    const size_t base = (size_t)frame * MAXB;
    BitImg im;
    im.bits = a.morph_bits + (size_t)frame * H * a.WW;
    im.W = W; im.H = H; im.WW = a.WW;
    const uint16_t* owner = a.owner + (size_t)frame * W * H;
    for (int b = tid; b < nb; b += FNT) { s_key[b] = a.key[base + b]; s_an[b] = a.an[base + b]; }
    __syncthreads();
    // hole -> outer border of its component
    const bool pieces = a.use_cuts && !meta->fallback;
    for (int b = tid; b < nb; b += FNT) {
        int cur = b;
        if (pieces) {
            while (s_key[cur] & 1u) cur = a.hole_owner[base + cur];
            s_parent[b] = (short)cur;
            continue;
        }
        while (s_key[cur] & 1u) {  // climb: left end of the foreground run left of the hole start
            int p = (int)(s_key[cur] >> 1), y = p / (W + 1), x = p - y * (W + 1);
            int xl = x - 1;
            while (im.at(xl - 1, y)) --xl;
            cur = owner[(size_t)y * W + xl];
        }
        s_parent[b] = (short)cur;
    }
    __syncthreads();
    // CCOMP output order: outers by descending rank, each followed by its holes, descending
    for (int b = tid; b < nb; b += FNT) {
        bool hole_b = s_key[b] & 1u;
        int top_b = s_parent[b];
        int pos = 0;
        for (int o = 0; o < nb; ++o) {
            if (o == b) continue;
            bool hole_o = s_key[o] & 1u;
            int top_o = s_parent[o];
            bool before;
            if (top_o != top_b) before = top_o > top_b;
            else if (hole_o != hole_b) before = !hole_o;
            else before = o > b;
            pos += before;
        }
        s_pos2b[pos] = (short)b;
    }
    __syncthreads();
    if (tid == 0) {
        int nraw = 0, qf = 0;
        for (int p = 0; p < nb; ++p) {
            int b = s_pos2b[p];
            if (s_an[b] != 4) continue;
            if (nraw >= MANTIS_MAX_QUADS) { qf = FRAME_FLAG_QUAD_OVERFLOW; break; }
            s_qb[nraw++] = (short)b;
        }
        s_nraw = nraw;
        s_qflag = qf;
    }
    __syncthreads();
    const int nraw = s_nraw;
    const short2* quad = a.quad + base * 4;
    for (int e = tid; e < nraw; e += FNT) {
        const short2* q = quad + 4 * s_qb[e];
        float xs = 0.f, ys = 0.f;
        for (int k = 0; k < 4; ++k) { xs += (float)q[k].x; ys += (float)q[k].y; }
        s_cx[e] = xs / 4.f;
        s_cy[e] = ys / 4.f;
        s_neighbor[e] = 0;
    }
    __syncthreads();
    if (tid < 32) {
        // removeDuplicateQuads: exact radius search, 4 results, unfilled slots alias quad 0.
        // Sequential over quads as in the reference (a quad marked by an earlier one is skipped); the
        // search of one quad is spread over the warp: every lane keeps (distance, index) keys of its
        // quads and the four smallest are taken one after the other (ties: the smaller index, which
        // is what the reference's stable insertion keeps).
        constexpr int PER = MANTIS_MAX_QUADS / 32;
        for (int e = 0; e < nraw; ++e) {
            if (s_neighbor[e]) continue;  // (uniform: read after the __syncwarp below)
            const short2* q = quad + 4 * s_qb[e];
            int ddx = q[0].x - q[1].x, ddy = q[0].y - q[1].y;
            double side = sqrt((double)(ddx * ddx + ddy * ddy));
            float radius = (float)(a.radius_mult * side);
            unsigned long long key[PER];
#pragma unroll
            for (int t = 0; t < PER; ++t) {
                const int j = tid + 32 * t;
                key[t] = ~0ull;
                if (j < nraw) {
                    float fx = s_cx[e] - s_cx[j], fy = s_cy[e] - s_cy[j];
                    float d = __fadd_rn(__fmul_rn(fx, fx), __fmul_rn(fy, fy));
                    if (d <= radius) key[t] = ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)j;
                }
            }
            int found = 0;
            for (; found < 4; ++found) {
                unsigned long long m = key[0];
#pragma unroll
                for (int t = 1; t < PER; ++t) m = key[t] < m ? key[t] : m;
                const unsigned dm = __reduce_min_sync(0xffffffffu, (unsigned)(m >> 32));
                if (dm == 0xffffffffu) break;
                const unsigned jm = __reduce_min_sync(0xffffffffu, (unsigned)(m >> 32) == dm ? (unsigned)m : 0xffffffffu);
                const unsigned long long taken = ((unsigned long long)dm << 32) | jm;
#pragma unroll
                for (int t = 0; t < PER; ++t)
                    if (key[t] == taken) key[t] = ~0ull;
                if (found >= 1 && tid == 0) s_neighbor[jm] = 1;
            }
            if (found < 4 && tid == 0) s_neighbor[0] = 1;
            __syncwarp();
        }
    }
    __syncthreads();
    if (tid == 0) {
        int nq = 0;
        for (int e = 0; e < nraw; ++e) {
            short slot = -1;
            if (!s_neighbor[e]) slot = (short)nq++;
            s_pos2b[e] = slot;  // reuse: raw quad -> output slot
        }
        out->n_contours = nb;
        out->flags = s_qflag;
        out->n_quads_raw = nraw;
        out->n_quads = nq;
    }
    __syncthreads();
    for (int e = tid; e < nraw; e += FNT) {
        const int nq = s_pos2b[e];
        if (nq < 0) continue;
        const short2* q = quad + 4 * s_qb[e];
        for (int k = 0; k < 4; ++k) {
            int px = q[k].x, py = q[k].y;
            out->corners[nq][2 * k] = px;
            out->corners[nq][2 * k + 1] = py;
            float ex = (float)px - s_cx[e], ey = (float)py - s_cy[e];
            float mx = (float)((double)ex * a.stretch), my = (float)((double)ey * a.stretch);
            out->test_px[nq][2 * k] = (double)__fadd_rn(mx, s_cx[e]);
            out->test_px[nq][2 * k + 1] = (double)__fadd_rn(my, s_cy[e]);
        }
    }
}

}  // namespace

int launch_quads(mantis_ctx* c, const uint32_t* d_morph_bits, int n, int w, int h, FrameQuads* d_out)
{
    int rc;
    const size_t nB = (size_t)n * MAXB;
    // one arena: meta | cand | ukey unp usteps key np off | bylen an | quad
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o += (bytes + 255) & ~(size_t)255; return at; };
    const size_t o_meta = take((size_t)n * sizeof(QMeta)), o_cand = take((size_t)n * MAXCAND * 4), o_ukey = take(nB * 4),
                 o_unp = take(nB * 4), o_ust = take(nB * 4), o_key = take(nB * 4), o_np = take(nB * 4), o_off = take(nB * 4),
                 o_bylen = take(nB * 2), o_long = take((size_t)n * MAXLONG * 2), o_an = take(nB * 2), o_quad = take(nB * 4 * sizeof(short2)),
                 o_cut = take((size_t)n * MAXCUT * 4), o_cend = take((size_t)n * MAXCUT * 4), o_cnp = take((size_t)n * MAXCUT * 4),
                 o_cdst = take((size_t)n * MAXCUT * 4), o_ckey = take((size_t)n * MAXCUT * 4),
                 o_cpar = take((size_t)n * MAXCUT * 4), o_hown = take(nB * 2);
    if ((rc = ctx_reserve(c, c->q_stack, o)) ||
        (rc = ctx_reserve(c, c->q_pts, (size_t)n * MANTIS_MAX_PTS * sizeof(short2) * 2)) ||
        (rc = ctx_reserve(c, c->q_dst, (size_t)n * (MANTIS_MAX_PTS + 2 * MAXB) * sizeof(int2))) ||
        (rc = ctx_reserve(c, c->q_owner, (size_t)n * w * h * sizeof(uint16_t))) ||
        (rc = ctx_reserve(c, c->q_tmp, (size_t)n * MAXCUT * TMPV * sizeof(short2))))
        return rc;
    char* ar = (char*)c->q_stack.p;
    QuadArgs a;
    a.morph_bits = d_morph_bits;
    a.W = w; a.H = h; a.WW = words_per_row(w); a.n_frames = n;
    a.meta = (QMeta*)(ar + o_meta); a.cand = (uint32_t*)(ar + o_cand); a.ukey = (uint32_t*)(ar + o_ukey);
    a.unp = (int*)(ar + o_unp); a.usteps = (int*)(ar + o_ust); a.key = (uint32_t*)(ar + o_key);
    a.np = (int*)(ar + o_np); a.off = (int*)(ar + o_off); a.bylen = (short*)(ar + o_bylen); a.longl = (short*)(ar + o_long);
    a.an = (short*)(ar + o_an); a.quad = (short2*)(ar + o_quad);
    a.cut = (uint32_t*)(ar + o_cut); a.cut_end = (uint32_t*)(ar + o_cend); a.cut_np = (int*)(ar + o_cnp); a.cut_dst = (uint32_t*)(ar + o_cdst);
    a.cut_key = (uint32_t*)(ar + o_ckey); a.cut_par = (uint32_t*)(ar + o_cpar); a.hole_owner = (short*)(ar + o_hown);
    a.pts = (short2*)c->q_pts.p;
    a.dst = a.pts + (size_t)n * MANTIS_MAX_PTS;
    a.stack = (int2*)c->q_dst.p;
    a.owner = (uint16_t*)c->q_owner.p;
    a.cut_tmp = (short2*)c->q_tmp.p;
    a.out = d_out;
    a.eps = (double)c->cfg.polygon_epsilon;
    a.stretch = c->cfg.quad_stretch_factor;
    a.radius_mult = c->cfg.search_radius_multiplier;
    MCHECK(c, cudaMemsetAsync(a.meta, 0, (size_t)n * sizeof(QMeta), c->stream));
    const size_t plane = (size_t)(h + 2) * (a.WW + 2) * sizeof(uint32_t);
    // 1 = the plane is staged in shared memory; 2 = it does not fit (1920x1080): zero-framed copy in global memory
    const int use_smem = plane <= 200 * 1024 ? 1 : 2;
    // piece keys hold (pixel << 3 | direction) in 24 bits
    a.use_cuts = (size_t)w * h < (1u << 21) && !getenv("MANTIS_QUADS_SEQUENTIAL");  // (the variable is for tests)
    a.cut_mask = (h <= 512 ? CUT_ROWS : h <= 1024 ? 2 * CUT_ROWS : 4 * CUT_ROWS) - 1;  // about the same number of cut rows per frame
    a.padded = nullptr; a.pad_words = plane / sizeof(uint32_t);
    if (use_smem == 1) {
        MCHECK(c, cudaFuncSetAttribute(k_qprobe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plane));
        MCHECK(c, cudaFuncSetAttribute(k_qtrace, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plane));
        MCHECK(c, cudaFuncSetAttribute(k_qseg<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plane));
        MCHECK(c, cudaFuncSetAttribute(k_qemit<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plane));
    } else {
        if ((rc = ctx_reserve(c, c->q_pad, (size_t)n * plane))) return rc;
        a.padded = (const uint32_t*)c->q_pad.p;
        k_qpad<<<dim3(8, n), CNT, 0, c->stream>>>(a, (uint32_t*)c->q_pad.p);
        c->launches += 1;
    }
    if (a.use_cuts) {
        k_qcut<<<dim3(4, n), CNT, 0, c->stream>>>(a);
        if (use_smem == 1) k_qseg<false><<<n, CNT, plane, c->stream>>>(a);
        else k_qseg<true><<<n, CNT, 0, c->stream>>>(a);
        k_qlink<<<n, LNT, 0, c->stream>>>(a);
        c->launches += 3;
    }
    // the sequential route: all frames when the plane does not fit in shared memory, else only the frames flagged above
    k_qcand<<<dim3(a.use_cuts ? 2 : 16, n), CNT, 0, c->stream>>>(a);
    k_qprobe<<<n, CNT, use_smem == 1 ? plane : 0, c->stream>>>(a, use_smem);
    k_qrank<<<n, CNT, 0, c->stream>>>(a);
    k_qtrace<<<n, TNT, use_smem == 1 ? plane : 0, c->stream>>>(a, use_smem);
    if (a.use_cuts) {
        k_qemit_copy<<<n, CNT, 0, c->stream>>>(a);
        if (use_smem == 1) k_qemit<false><<<n, CNT, plane, c->stream>>>(a);
        else k_qemit<true><<<n, CNT, 0, c->stream>>>(a);
        c->launches += 2;
    }
    k_qapprox_long<<<n, 32 * LWARPS, 0, c->stream>>>(a);
    k_qfinish<<<n, FNT, 0, c->stream>>>(a);
    c->launches += 6;
    MCHECK(c, cudaGetLastError());
    if (getenv("MANTIS_QUAD_DEBUG")) {  // how many frames took which route
        std::vector<QMeta> hm(n);
        MCHECK(c, cudaMemcpyAsync(hm.data(), a.meta, (size_t)n * sizeof(QMeta), cudaMemcpyDeviceToHost, c->stream));
        MCHECK(c, cudaStreamSynchronize(c->stream));
        long long fb = 0, cuts = 0, maxc = 0;
        for (const QMeta& m : hm) { fb += m.fallback != 0; cuts += m.ncut; maxc = m.ncut > maxc ? m.ncut : maxc; }
        fprintf(stderr, "quads: %d frames, %lld sequential, cuts mean %.1f max %lld\n", n, fb, (double)cuts / (n ? n : 1), maxc);
    }
    return MANTIS_OK;
}
