/*
 * approx.c -- oracle (test infrastructure) for cv::approxPolyDP(curve, eps, closed)
 * as called at include/mantis3/QuadDetection.h:221 (eps = POLYGON_EPSILON, closed = true).
 * Restates OpenCV's Douglas-Peucker variant (two-farthest-points initialisation for
 * closed curves, explicit slice stack, final collinear clean-up); pinned against
 * cv2.approxPolyDP by tests/test_oracle_contours.py.
 */
#include "mantis_oracle.h"
#include <math.h>
#include <stdlib.h>

typedef struct { int start, end; } range_t;
typedef struct { int x, y; } ipt;

int mo_approx_poly_dp(const int* src_xy, int count0, double eps, int closed0, int* dst_xy)
{
    const ipt* src = (const ipt*)src_xy;
    ipt* dst = (ipt*)dst_xy;
    int init_iters = 3;
    range_t slice = {0, 0}, right_slice = {0, 0};
    ipt start_pt = {-1000000, -1000000}, end_pt = {0, 0}, pt = {0, 0};
    int i = 0, j, pos = 0, wpos, count = count0, new_count = 0;
    int is_closed = closed0;
    int le_eps = 0;
    size_t top = 0;
    if (count == 0) return 0;
    range_t* stack = (range_t*)malloc(sizeof(range_t) * (size_t)(count + 8) * 2);

#define READ_PT(p_, pos_) do { p_ = src[pos_]; if (++pos_ >= count) pos_ = 0; } while (0)
#define READ_DST_PT(p_, pos_) do { p_ = dst[pos_]; if (++pos_ >= count) pos_ = 0; } while (0)
#define WRITE_PT(p_) dst[new_count++] = p_
#define PUSH_SLICE(s_) stack[top++] = s_

    eps *= eps;
    if (!is_closed) {
        right_slice.start = count;
        end_pt = src[0];
        start_pt = src[count - 1];
        if (start_pt.x != end_pt.x || start_pt.y != end_pt.y) {
            slice.start = 0;
            slice.end = count - 1;
            PUSH_SLICE(slice);
        } else {
            is_closed = 1;
            init_iters = 1;
        }
    }
    if (is_closed) {
        /* 1. find approximately two farthest points of the contour */
        right_slice.start = 0;
        for (i = 0; i < init_iters; i++) {
            double dist, max_dist = 0;
            pos = (pos + right_slice.start) % count;
            READ_PT(start_pt, pos);
            for (j = 1; j < count; j++) {
                double dx, dy;
                READ_PT(pt, pos);
                dx = pt.x - start_pt.x;
                dy = pt.y - start_pt.y;
                dist = dx * dx + dy * dy;
                if (dist > max_dist) {
                    max_dist = dist;
                    right_slice.start = j;
                }
            }
            le_eps = max_dist <= eps;
        }
        /* 2. initialise the stack */
        if (!le_eps) {
            right_slice.end = slice.start = pos % count;
            slice.end = right_slice.start = (right_slice.start + slice.start) % count;
            PUSH_SLICE(right_slice);
            PUSH_SLICE(slice);
        } else {
            WRITE_PT(start_pt);
        }
    }
    /* 3. recursive subdivision */
    while (top > 0) {
        slice = stack[--top];
        end_pt = src[slice.end];
        pos = slice.start;
        READ_PT(start_pt, pos);
        if (pos != slice.end) {
            double dx, dy, dist, max_dist = 0;
            dx = end_pt.x - start_pt.x;
            dy = end_pt.y - start_pt.y;
            /* OpenCV 4.13 measures the distance to the SEGMENT start-end (not the infinite
             * line): beyond either end it is the distance to that end point.  All terms are
             * kept as exact integers scaled by L = |end-start|^2. */
            const double L = dx * dx + dy * dy;
            while (pos != slice.end) {
                READ_PT(pt, pos);
                double px = pt.x - start_pt.x, py = pt.y - start_pt.y;
                double t = px * dx + py * dy;
                if (t <= 0) {
                    dist = (px * px + py * py) * L;
                } else if (t >= L) {
                    double qx = pt.x - end_pt.x, qy = pt.y - end_pt.y;
                    dist = (qx * qx + qy * qy) * L;
                } else {
                    double cr = py * dx - px * dy;
                    dist = cr * cr;
                }
                if (dist > max_dist) {
                    max_dist = dist;
                    right_slice.start = (pos + count - 1) % count;
                }
            }
            le_eps = max_dist <= eps * L;
        } else {
            le_eps = 1;
            start_pt = src[slice.start];
        }
        if (le_eps) {
            WRITE_PT(start_pt);
        } else {
            right_slice.end = slice.end;
            slice.end = right_slice.start;
            PUSH_SLICE(right_slice);
            PUSH_SLICE(slice);
        }
    }
    /* 4. clean-up: drop points on (almost) straight lines */
    is_closed = closed0;
    count = new_count;
    pos = is_closed ? count - 1 : 0;
    READ_DST_PT(start_pt, pos);
    wpos = pos;
    READ_DST_PT(pt, pos);
    for (i = !is_closed; i < count - !is_closed && new_count > 2; i++) {
        double dx, dy, dist, sip;
        READ_DST_PT(end_pt, pos);
        dx = end_pt.x - start_pt.x;
        dy = end_pt.y - start_pt.y;
        dist = fabs((pt.x - start_pt.x) * dy - (pt.y - start_pt.y) * dx);
        sip = (double)(pt.x - start_pt.x) * (end_pt.x - pt.x) + (double)(pt.y - start_pt.y) * (end_pt.y - pt.y);
        if (dist * dist <= 0.5 * eps * (dx * dx + dy * dy) && dx != 0 && dy != 0 && sip >= 0) {
            new_count--;
            dst[wpos] = start_pt = end_pt;
            if (++wpos >= count) wpos = 0;
            READ_DST_PT(pt, pos);
            i++;
            continue;
        }
        dst[wpos] = start_pt = pt;
        if (++wpos >= count) wpos = 0;
        pt = end_pt;
    }
    if (!is_closed) dst[wpos] = pt;
    free(stack);
    return new_count;
}
