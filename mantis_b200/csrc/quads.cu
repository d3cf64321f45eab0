// quads.cu -- K5: grid-cell (quad) extraction from the bit-packed morphology image.
//
// Realises, per frame and with OpenCV's exact results AND output order,
//   cv::findContours(RETR_CCOMP, CHAIN_APPROX_SIMPLE)       include/mantis3/QuadDetection.h:216
//   cv::approxPolyDP(contour, POLYGON_EPSILON, closed)      :221   (keep 4-gons, :222-225)
//   Quadrilateral ctor (centre, side, stretched test points)  :13-66
//   removeDuplicateQuads                                     :115-171
//
// Suzuki-Abe border following is a sequential raster scan in OpenCV; here every border is found
// and traced independently:
//   1. candidates: a border can only start at a local top-left extreme -- an outer border at a
//      foreground pixel with no foreground W/NW/N/NE neighbour, a hole border at a background
//      pixel whose W and N neighbours are foreground (bit tests on whole words).
//   2. one thread per candidate follows its border and gives up as soon as it meets a pixel that
//      the raster scan would have reached earlier (or, for holes, the image frame); a candidate
//      that closes its loop is exactly a border start of the sequential algorithm.
//   3. starts are ranked in raster order (= OpenCV's discovery order); each border is traced
//      again to emit its CHAIN_APPROX_SIMPLE points and to record which border owns each
//      left edge it passes (needed for the hole -> component relation of RETR_CCOMP).
//   4. approxPolyDP per border (OpenCV 4.13 variant: distance to the SEGMENT), CCOMP ordering
//      (outer borders in reverse discovery order, each followed by its holes, reversed).
#include "ctx.h"
#include "pipeline.cuh"

namespace {

constexpr int QNT = 256;
constexpr int MAXB = MANTIS_MAX_BORDERS;   // border starts per frame
constexpr int MAXCAND = 4096;              // queued candidates (more are walked inline)

struct BitImg {
    const uint32_t* bits;
    int W, H, WW;
    __device__ __forceinline__ int at(int x, int y) const
    {
        if ((unsigned)x >= (unsigned)W || (unsigned)y >= (unsigned)H) return 0;
        return (__ldg(bits + (size_t)y * WW + (x >> 5)) >> (x & 31)) & 1;
    }
};

__constant__ int c_dx[8] = {1, 1, 0, -1, -1, -1, 0, 1};
__constant__ int c_dy[8] = {0, -1, -1, -1, 0, 1, 1, 1};

// Border following (OpenCV's icvFetchContour on a 0/1 image with a zero frame).
// Visitor:  bool step(x, y, zmask)  -- called for every border pixel with the set of directions
//           examined as background from it; return false to abort.
//           void point(x, y)        -- CHAIN_APPROX_SIMPLE vertex.
// Returns true if the loop closed.
template <class V>
__device__ bool follow_border(const BitImg& im, int x0, int y0, bool is_hole, V& vis)
{
    int s_end = is_hole ? 0 : 4, s = s_end;
    int nz;
    do {
        s = (s - 1) & 7;
        nz = im.at(x0 + c_dx[s], y0 + c_dy[s]);
    } while (!nz && s != s_end);
    if (s == s_end) {  // isolated pixel
        if (!vis.step(x0, y0, 0xffu)) return false;
        vis.point(x0, y0);
        return true;
    }
    const int x1 = x0 + c_dx[s], y1 = y0 + c_dy[s];
    int x3 = x0, y3 = y0, prev_s = s ^ 4;
    for (;;) {
        s_end = s;
        unsigned zmask = 0;
        int x4, y4;
        for (;;) {
            ++s;
            x4 = x3 + c_dx[s & 7];
            y4 = y3 + c_dy[s & 7];
            if (im.at(x4, y4)) break;
            zmask |= 1u << (s & 7);
        }
        s &= 7;
        if (!vis.step(x3, y3, zmask)) return false;
        if (s != prev_s) {
            vis.point(x3, y3);
            prev_s = s;
        }
        if (x4 == x0 && y4 == y0 && x3 == x1 && y3 == y1) break;
        x3 = x4;
        y3 = y4;
        s = (s + 4) & 7;
    }
    return true;
}

// detection visitor: outer borders
struct OuterProbe {
    int x0, y0, npts;
    __device__ bool step(int x, int y, unsigned) { return !(y < y0 || (y == y0 && x < x0)); }
    __device__ void point(int, int) { ++npts; }
};
// detection visitor: hole borders; (hx, hy) is the hole's first background pixel
struct HoleProbe {
    int hx, hy, W, H, npts;
    __device__ bool step(int x, int y, unsigned zmask)
    {
#pragma unroll
        for (int d = 0; d < 8; d += 2) {
            if (!(zmask & (1u << d))) continue;
            int bx = x + c_dx[d], by = y + c_dy[d];
            if ((unsigned)bx >= (unsigned)W || (unsigned)by >= (unsigned)H) return false;  // touches the frame
            if (by < hy || (by == hy && bx < hx)) return false;
        }
        return true;
    }
    __device__ void point(int, int) { ++npts; }
};
// extraction visitor
struct Emit {
    short2* out;
    uint16_t* owner;
    int W, n;
    uint16_t id;
    __device__ bool step(int x, int y, unsigned zmask)
    {
        if (zmask & (1u << 4)) owner[(size_t)y * W + x] = id;  // this border owns the left edge of (x, y)
        return true;
    }
    __device__ void point(int x, int y) { out[n++] = make_short2((short)x, (short)y); }
};

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
    // clean-up of (almost) collinear vertices
    count = new_count;
    pos = count - 1;
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

}  // namespace

#include "quads_kernel.inl"
