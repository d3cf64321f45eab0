/*
 * hough.c -- oracle (TEST INFRASTRUCTURE ONLY, see mantis_oracle.h) for row a6' of SURVEY.md 8(a):
 * the north star's "Hough / line-voting accumulator, line clustering, grid-intersection extraction".
 *
 * The reference (k-sheridan/mantis) has NO Hough transform; its grid primitives are cell quads
 * (include/mantis3/QuadDetection.h:216-226).  The voting + peak stage therefore restates the published
 * algorithm of OpenCV's cv::HoughLines (HoughLinesStandard, modules/imgproc/src/hough.cpp of the 4.x
 * series, the third-party library every other stage of the reference calls) and is pinned against
 * cv2 4.13.0 (tests/golden/hough_golden.json, made by tests/golden/make_golden_hough.py).  Line
 * clustering and intersections have no third-party counterpart either: they are defined here, and the
 * CUDA implementation is checked against this file only ("parity unpinned by the reference").
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "mantis_oracle.h"

#define MO_PI 3.1415926535897932384626433832795

typedef struct mo_hough_line { float rho, theta; int votes; } mo_hough_line;

static int hough_numangle(double min_theta, double max_theta, double theta_step)
{
    int numangle = (int)floor((max_theta - min_theta) / theta_step) + 1;
    if (numangle > 1 && fabs(MO_PI - (numangle - 1) * theta_step) < theta_step / 2) --numangle;
    return numangle;
}

/* accumulator geometry shared with the CUDA side */
void mo_hough_dims(int w, int h, double rho, double theta, int* numangle, int* numrho)
{
    int max_rho = w + h, min_rho = -max_rho;
    *numangle = hough_numangle(0.0, MO_PI, theta);
    *numrho = (int)lrint(((max_rho - min_rho) + 1) / rho);
}

void mo_hough_trig(int numangle, double theta, double rho, float* tab_sin, float* tab_cos)
{
    float irho = 1 / (float)rho;
    float ang = 0.f;
    for (int n = 0; n < numangle; ang += (float)theta, n++) {
        tab_sin[n] = (float)(sin((double)ang) * irho);
        tab_cos[n] = (float)(cos((double)ang) * irho);
    }
}

typedef struct { const int* accum; } cmp_ctx;
static const int* g_accum;
static int cmp_gt(const void* a, const void* b)
{
    int l1 = *(const int*)a, l2 = *(const int*)b;
    if (g_accum[l1] != g_accum[l2]) return g_accum[l1] > g_accum[l2] ? -1 : 1;
    return l1 < l2 ? -1 : (l1 > l2 ? 1 : 0);
}

/* cv::HoughLines(edges, rho, theta, threshold) with min_theta 0, max_theta pi; returns the number of
 * lines written (at most max_lines), strongest first.  accum_out (optional) receives the
 * (numangle+2) x (numrho+2) accumulator. */
int mo_hough_lines(const uint8_t* edges, int w, int h, double rho, double theta, int threshold, mo_hough_line* lines,
                   int max_lines, int* accum_out)
{
    int numangle, numrho;
    mo_hough_dims(w, h, rho, theta, &numangle, &numrho);
    float* tab_sin = (float*)malloc(sizeof(float) * numangle);
    float* tab_cos = (float*)malloc(sizeof(float) * numangle);
    mo_hough_trig(numangle, theta, rho, tab_sin, tab_cos);
    size_t cells = (size_t)(numangle + 2) * (numrho + 2);
    int* accum = (int*)calloc(cells, sizeof(int));
    for (int i = 0; i < h; i++)
        for (int j = 0; j < w; j++)
            if (edges[(size_t)i * w + j] != 0)
                for (int n = 0; n < numangle; n++) {
                    volatile float a = j * tab_cos[n], b = i * tab_sin[n]; /* no contraction: two products, one sum */
                    int r = (int)lrintf(a + b);
                    r += (numrho - 1) / 2;
                    accum[(n + 1) * (numrho + 2) + r + 1]++;
                }
    int* sort_buf = (int*)malloc(sizeof(int) * (size_t)numangle * numrho);
    int total = 0;
    for (int r = 0; r < numrho; r++)
        for (int n = 0; n < numangle; n++) {
            int base = (n + 1) * (numrho + 2) + r + 1;
            if (accum[base] > threshold && accum[base] > accum[base - 1] && accum[base] >= accum[base + 1] &&
                accum[base] > accum[base - numrho - 2] && accum[base] >= accum[base + numrho + 2])
                sort_buf[total++] = base;
        }
    g_accum = accum;
    qsort(sort_buf, total, sizeof(int), cmp_gt);
    int n_out = total < max_lines ? total : max_lines;
    double scale = 1. / (numrho + 2);
    for (int i = 0; i < n_out; i++) {
        int idx = sort_buf[i];
        int n = (int)floor(idx * scale) - 1;
        int r = idx - (n + 1) * (numrho + 2) - 1;
        lines[i].rho = (r - (numrho - 1) * 0.5f) * (float)rho;
        lines[i].theta = (float)(0.0) + n * (float)theta;
        lines[i].votes = accum[idx];
    }
    if (accum_out) memcpy(accum_out, accum, cells * sizeof(int));
    free(sort_buf); free(accum); free(tab_sin); free(tab_cos);
    return n_out;
}

/* ---- line clustering and grid intersections (defined by this repo; no third-party counterpart) ----
 *
 * Lines arrive strongest first.  Greedy clustering: a line joins the first earlier seed it is similar
 * to, else it becomes a seed.  Similar: |d theta| <= theta_tol and |d rho| <= rho_tol, where a line
 * more than pi/2 away in angle is first re-expressed as (-rho, theta -+ pi).  A cluster's line is the
 * vote-weighted mean of its members, accumulated in integers (2*rho and the angle index are integers
 * for rho = 1), so the result does not depend on summation order. */
typedef struct mo_hough_cluster {
    float rho, theta;
    int votes, members;
} mo_hough_cluster;

int mo_hough_cluster_lines(const mo_hough_line* lines, int n, double theta_step, int numangle, double rho_tol,
                           double theta_tol, mo_hough_cluster* out, int max_clusters)
{
    int* seed_of = (int*)malloc(sizeof(int) * (n > 0 ? n : 1));
    int nc = 0;
    long long* s_v = (long long*)calloc(max_clusters, sizeof(long long));
    long long* s_r = (long long*)calloc(max_clusters, sizeof(long long));
    long long* s_a = (long long*)calloc(max_clusters, sizeof(long long));
    int* seed_idx = (int*)malloc(sizeof(int) * max_clusters);
    int* members = (int*)calloc(max_clusters, sizeof(int));
    for (int i = 0; i < n; i++) {
        int ai = (int)lrint(lines[i].theta / (float)theta_step);  /* angle index */
        long long r2 = (long long)lrint(2.0 * lines[i].rho);
        int c_found = -1, a_use = ai;
        long long r_use = r2;
        for (int c = 0; c < nc; c++) {
            const mo_hough_line* s = &lines[seed_idx[c]];
            int as = (int)lrint(s->theta / (float)theta_step);
            long long rs2 = (long long)lrint(2.0 * s->rho);
            int a = ai;
            long long r = r2;
            if (abs(a - as) * 2 > numangle) { a = a > as ? a - numangle : a + numangle; r = -r; }
            if (fabs((double)(a - as)) * theta_step <= theta_tol && fabs((double)(r - rs2)) * 0.5 <= rho_tol) {
                c_found = c; a_use = a; r_use = r;
                break;
            }
        }
        if (c_found < 0) {
            if (nc >= max_clusters) { seed_of[i] = -1; continue; }
            c_found = nc++;
            seed_idx[c_found] = i;
        }
        seed_of[i] = c_found;
        s_v[c_found] += lines[i].votes;
        s_r[c_found] += (long long)lines[i].votes * r_use;
        s_a[c_found] += (long long)lines[i].votes * a_use;
        members[c_found]++;
    }
    for (int c = 0; c < nc; c++) {
        out[c].rho = (float)((double)s_r[c] / (2.0 * (double)s_v[c]));
        out[c].theta = (float)(((double)s_a[c] / (double)s_v[c]) * theta_step);
        out[c].votes = (int)s_v[c];
        out[c].members = members[c];
    }
    free(seed_of); free(s_v); free(s_r); free(s_a); free(seed_idx); free(members);
    return nc;
}

/* Two dominant pencils: A = clusters within pencil_tol (mod pi) of the strongest cluster's angle,
 * B = clusters within pencil_tol of the strongest cluster that is at least pi/4 away from A's.
 * Intersections of every (a in A, b in B) pair, a-major in cluster order, kept when inside the
 * image.  xy receives (x, y) float pairs; pencil (optional, n entries) receives 0 = A, 1 = B, -1. */
static double ang_dist(double a, double b)
{
    double d = fabs(a - b);
    d = fmod(d, MO_PI);
    return d > MO_PI / 2 ? MO_PI - d : d;
}

int mo_hough_intersections(const mo_hough_cluster* cl, int n, int w, int h, double pencil_tol, float* xy,
                           int max_points, int* pencil)
{
    int* pen = (int*)malloc(sizeof(int) * (n > 0 ? n : 1));
    for (int i = 0; i < n; i++) pen[i] = -1;
    int count = 0;
    if (n > 0) {
        int sa = 0, sb = -1;
        for (int i = 1; i < n; i++)
            if (cl[i].votes > cl[sa].votes) sa = i;
        for (int i = 0; i < n; i++)
            if (ang_dist(cl[i].theta, cl[sa].theta) >= MO_PI / 4 && (sb < 0 || cl[i].votes > cl[sb].votes)) sb = i;
        for (int i = 0; i < n; i++) {
            if (ang_dist(cl[i].theta, cl[sa].theta) <= pencil_tol) pen[i] = 0;
            else if (sb >= 0 && ang_dist(cl[i].theta, cl[sb].theta) <= pencil_tol) pen[i] = 1;
        }
        for (int a = 0; a < n; a++) {
            if (pen[a] != 0) continue;
            double ca = cos((double)cl[a].theta), sna = sin((double)cl[a].theta);
            for (int b = 0; b < n; b++) {
                if (pen[b] != 1) continue;
                double cb = cos((double)cl[b].theta), snb = sin((double)cl[b].theta);
                double det = ca * snb - sna * cb;
                if (fabs(det) < 1e-9) continue;
                double x = ((double)cl[a].rho * snb - (double)cl[b].rho * sna) / det;
                double y = ((double)cl[b].rho * ca - (double)cl[a].rho * cb) / det;
                if (!(x >= 0 && x < w && y >= 0 && y < h)) continue;
                if (count < max_points) { xy[2 * count] = (float)x; xy[2 * count + 1] = (float)y; }
                count++;
            }
        }
    }
    if (pencil) memcpy(pencil, pen, sizeof(int) * n);
    free(pen);
    return count < max_points ? count : max_points;
}
