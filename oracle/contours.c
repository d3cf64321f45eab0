/*
 * contours.c -- oracle (test infrastructure, see mantis_oracle.h) for
 *   cv::findContours(RETR_CCOMP | RETR_LIST, CHAIN_APPROX_SIMPLE)   QuadDetection.h:216,
 *                                                                   HypothesisEvaluation.h:343
 *   cv::approxPolyDP(contour, eps, closed=true)                     QuadDetection.h:221
 *
 * OpenCV is not vendored by the reference; this restates the published algorithm
 * (Suzuki & Abe 1985 border following as implemented by OpenCV, and OpenCV's
 * Douglas-Peucker variant for closed curves) and is pinned against cv2 4.13.0 on random
 * and fixture images by tests/test_oracle_contours.py (point lists AND output order).
 */
#include "mantis_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int* pts;      /* x,y pairs of all contours, concatenated */
    int npts, cap_pts;
    mo_contour* c; /* per-contour records in discovery order */
    int nc, cap_c;
} builder;

static void push_pt(builder* b, int x, int y)
{
    if (b->npts == b->cap_pts) {
        b->cap_pts = b->cap_pts ? 2 * b->cap_pts : 4096;
        b->pts = (int*)realloc(b->pts, sizeof(int) * 2 * (size_t)b->cap_pts);
    }
    b->pts[2 * b->npts] = x;
    b->pts[2 * b->npts + 1] = y;
    b->npts++;
}

/* 8 directions, counter-clockwise starting at +x: E, NE, N, NW, W, SW, S, SE */
static const int DX[8] = {1, 1, 0, -1, -1, -1, 0, 1};
static const int DY[8] = {0, -1, -1, -1, 0, 1, 1, 1};

/* Follow one border starting at padded position (x, y).  img: int labels with a 1-px
 * zero frame, step = padded width.  Marks pixels with nbd (or -nbd on right edges). */
static void follow_border(int* img, int step, int x, int y, int is_hole, int nbd, builder* b)
{
    int delta[16];
    for (int k = 0; k < 8; ++k) delta[k] = delta[k + 8] = DY[k] * step + DX[k];
    int* i0 = img + (size_t)y * step + x;
    int* i1;
    int* i3;
    int* i4;
    int s_end, s, prev_s;
    int px = x - 1, py = y - 1; /* report unpadded coordinates */

    s_end = s = is_hole ? 0 : 4;
    do {
        s = (s - 1) & 7;
        i1 = i0 + delta[s];
    } while (*i1 == 0 && s != s_end);

    if (s == s_end) { /* isolated pixel */
        *i0 = -nbd;
        push_pt(b, px, py);
        return;
    }
    i3 = i0;
    prev_s = s ^ 4;
    for (;;) {
        s_end = s;
        for (;;) {
            i4 = i3 + delta[++s];
            if (*i4 != 0) break;
        }
        s &= 7;
        if ((unsigned)(s - 1) < (unsigned)s_end)
            *i3 = -nbd; /* the zero pixel to the right was examined: right edge */
        else if (*i3 == 1)
            *i3 = nbd;
        if (s != prev_s) {
            push_pt(b, px, py);
            prev_s = s;
        }
        px += DX[s];
        py += DY[s];
        if (i4 == i0 && i3 == i1) break;
        i3 = i4;
        s = (s + 4) & 7;
    }
}

/*
 * Raster scan.  Contours are recorded in discovery order with:
 *   is_hole, parent (index of the border that encloses it as in Suzuki's parent rule,
 *   -1 = frame), start point.  Output ordering (LIST / CCOMP) is applied afterwards.
 */
int mo_find_contours(const uint8_t* src, int w, int h, int mode, mo_contours* out)
{
    int step = w + 2;
    int* img = (int*)calloc((size_t)step * (h + 2), sizeof(int));
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) img[(size_t)(y + 1) * step + x + 1] = src[(size_t)y * w + x] ? 1 : 0;

    builder b;
    memset(&b, 0, sizeof(b));
    int nbd = 1;
    for (int y = 1; y <= h; ++y) {
        int* row = img + (size_t)y * step;
        int prev = 0;
        int lnbd = 0; /* label of the last labelled border pixel met on this row (0 = frame) */
        for (int x = 1; x <= w + 1; ++x) {
            int p = row[x];
            if (p == prev) continue;
            int is_hole = 0, start = 0;
            if (prev == 0 && p == 1) {
                start = 1;
            } else if (p == 0 && prev >= 1) {
                if (prev > 1) lnbd = prev;
                is_hole = 1;
                start = 1;
            }
            if (start) {
                ++nbd; /* labels start at 2 */
                if (b.nc == b.cap_c) {
                    b.cap_c = b.cap_c ? 2 * b.cap_c : 256;
                    b.c = (mo_contour*)realloc(b.c, sizeof(mo_contour) * (size_t)b.cap_c);
                }
                mo_contour* c = &b.c[b.nc];
                c->is_hole = is_hole;
                c->first = b.npts;
                /* Suzuki's parent rule from the last border met (index = label - 2) */
                int parent = -1;
                if (lnbd >= 2) {
                    int li = lnbd - 2;
                    parent = (b.c[li].is_hole == is_hole) ? b.c[li].parent : li;
                }
                c->parent = parent;
                follow_border(img, step, x - is_hole, y, is_hole, nbd, &b);
                c = &b.c[b.nc];
                c->count = b.npts - c->first;
                b.nc++;
                p = row[x];
            }
            prev = p;
            if (abs(p) > 1) lnbd = abs(p);
        }
    }
    free(img);

    /* output order */
    int* order = (int*)malloc(sizeof(int) * (size_t)(b.nc ? b.nc : 1));
    int no = 0;
    if (mode == MO_RETR_LIST) {
        for (int i = b.nc - 1; i >= 0; --i) order[no++] = i;
    } else { /* CCOMP: outer borders in reverse discovery order, each followed by its holes, reversed */
        for (int i = b.nc - 1; i >= 0; --i) {
            if (b.c[i].is_hole) continue;
            order[no++] = i;
            for (int j = b.nc - 1; j >= 0; --j)
                if (b.c[j].is_hole && b.c[j].parent == i) order[no++] = j;
        }
    }
    out->n = no;
    out->c = (mo_contour*)malloc(sizeof(mo_contour) * (size_t)(no ? no : 1));
    out->pts = (int*)malloc(sizeof(int) * 2 * (size_t)(b.npts ? b.npts : 1));
    int w_ = 0;
    for (int k = 0; k < no; ++k) {
        mo_contour c = b.c[order[k]];
        memcpy(out->pts + 2 * (size_t)w_, b.pts + 2 * (size_t)c.first, sizeof(int) * 2 * (size_t)c.count);
        c.first = w_;
        c.discovery = order[k];
        out->c[k] = c;
        w_ += c.count;
    }
    out->npts = w_;
    free(order);
    free(b.pts);
    free(b.c);
    return no;
}

void mo_free_contours(mo_contours* c)
{
    free(c->c);
    free(c->pts);
    c->c = NULL;
    c->pts = NULL;
    c->n = c->npts = 0;
}
