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
