/*
 * image_stage.c -- oracle (test infrastructure, see mantis_oracle.h) for the
 * image stage of detectQuadrilaterals(), include/mantis3/QuadDetection.h:203-214:
 *
 *   cv::cvtColor(BGR2GRAY)                 :209
 *   cv::GaussianBlur(3x3, sigma 3)         :211
 *   cv::Canny(T, 3T, aperture 3, L1)       :212
 *   cv::dilate(iter 2); cv::erode(iter 1)  :213-214
 *
 * OpenCV is a third-party dependency of the reference and is not vendored in its
 * tree; the arithmetic below restates OpenCV 4.13.0's 8-bit paths and is pinned
 * bit-for-bit against cv2 4.13.0 by tests/test_oracle_image.py.
 */
#include "mantis_oracle.h"
#include <stdlib.h>
#include <string.h>

/* cvtColor BGR2GRAY, 8-bit: Q15 weights B 3735, G 19235, R 9798, round-to-nearest. */
void mo_bgr2gray(const uint8_t* bgr, int w, int h, int stride, uint8_t* gray)
{
    for (int y = 0; y < h; ++y) {
        const uint8_t* p = bgr + (size_t)y * stride;
        uint8_t* g = gray + (size_t)y * w;
        for (int x = 0; x < w; ++x, p += 3)
            g[x] = (uint8_t)((3735 * p[0] + 19235 * p[1] + 9798 * p[2] + 16384) >> 15);
    }
}

static inline int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        else i = 2 * n - 2 - i;
    }
    return i;
}

/* GaussianBlur 3x3 sigma=3, 8-bit: fixed-point taps (84, 88, 84)/256 per axis,
 * BORDER_REFLECT_101, one rounding at the end: (v + 2^15) >> 16. */
void mo_blur3(const uint8_t* src, int w, int h, uint8_t* dst)
{
    uint32_t* hrow = (uint32_t*)malloc((size_t)w * h * sizeof(uint32_t));
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = src + (size_t)y * w;
        uint32_t* o = hrow + (size_t)y * w;
        for (int x = 0; x < w; ++x) {
            int xl = reflect101(x - 1, w), xr = reflect101(x + 1, w);
            o[x] = 84u * s[xl] + 88u * s[x] + 84u * s[xr];
        }
    }
    for (int y = 0; y < h; ++y) {
        const uint32_t* a = hrow + (size_t)reflect101(y - 1, h) * w;
        const uint32_t* b = hrow + (size_t)y * w;
        const uint32_t* c = hrow + (size_t)reflect101(y + 1, h) * w;
        uint8_t* o = dst + (size_t)y * w;
        for (int x = 0; x < w; ++x)
            o[x] = (uint8_t)((84u * a[x] + 88u * b[x] + 84u * c[x] + 32768u) >> 16);
    }
    free(hrow);
}

static inline int clampi(int i, int n) { return i < 0 ? 0 : (i >= n ? n - 1 : i); }

/* Sobel 3x3, CV_16S, scale 1, BORDER_REPLICATE (what cv::Canny asks for). */
void mo_sobel3(const uint8_t* src, int w, int h, int16_t* dx, int16_t* dy)
{
    for (int y = 0; y < h; ++y) {
        const uint8_t* r0 = src + (size_t)clampi(y - 1, h) * w;
        const uint8_t* r1 = src + (size_t)y * w;
        const uint8_t* r2 = src + (size_t)clampi(y + 1, h) * w;
        for (int x = 0; x < w; ++x) {
            int xl = clampi(x - 1, w), xr = clampi(x + 1, w);
            int gx = (r0[xr] - r0[xl]) + 2 * (r1[xr] - r1[xl]) + (r2[xr] - r2[xl]);
            int gy = (r2[xl] + 2 * r2[x] + r2[xr]) - (r0[xl] + 2 * r0[x] + r0[xr]);
            dx[(size_t)y * w + x] = (int16_t)gx;
            dy[(size_t)y * w + x] = (int16_t)gy;
        }
    }
}

/* cv::Canny(src, dst, low, high, 3, L2gradient=false): L1 magnitude with a zero
 * border, fixed-point direction test (TG22 = 13573 in Q15), double threshold,
 * stack-driven 8-connected hysteresis. */
void mo_canny(const uint8_t* src, int w, int h, int low, int high, uint8_t* dst)
{
    size_t n = (size_t)w * h;
    int16_t* dx = (int16_t*)malloc(n * sizeof(int16_t));
    int16_t* dy = (int16_t*)malloc(n * sizeof(int16_t));
    mo_sobel3(src, w, h, dx, dy);

    int ms = w + 2; /* magnitude / map buffers carry a 1-px border */
    int* mag = (int*)calloc((size_t)ms * (h + 2), sizeof(int));
    /* map: 1 = not an edge, 0 = candidate (weak), 2 = edge */
    uint8_t* map = (uint8_t*)malloc((size_t)ms * (h + 2));
    memset(map, 1, (size_t)ms * (h + 2));
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            int gx = dx[(size_t)y * w + x], gy = dy[(size_t)y * w + x];
            mag[(size_t)(y + 1) * ms + x + 1] = abs(gx) + abs(gy);
        }

    size_t* stack = (size_t*)malloc(n * sizeof(size_t));
    size_t sp = 0;
    const int TG22 = 13573;
    for (int y = 0; y < h; ++y) {
        const int* mp = mag + (size_t)y * ms + 1;       /* previous row */
        const int* ma = mag + (size_t)(y + 1) * ms + 1; /* this row     */
        const int* mn = mag + (size_t)(y + 2) * ms + 1; /* next row     */
        uint8_t* mrow = map + (size_t)(y + 1) * ms + 1;
        for (int x = 0; x < w; ++x) {
            int m = ma[x];
            int is_edge = 0;
            if (m > low) {
                int xs = dx[(size_t)y * w + x], ys = dy[(size_t)y * w + x];
                int ax = abs(xs), ay = abs(ys) << 15;
                int tg22x = ax * TG22;
                if (ay < tg22x) {
                    is_edge = (m > ma[x - 1] && m >= ma[x + 1]);
                } else {
                    int tg67x = tg22x + (ax << 16);
                    if (ay > tg67x) {
                        is_edge = (m > mp[x] && m >= mn[x]);
                    } else {
                        int s = ((xs ^ ys) < 0) ? -1 : 1;
                        is_edge = (m > mp[x - s] && m > mn[x + s]);
                    }
                }
            }
            if (is_edge) {
                if (m > high) {
                    mrow[x] = 2;
                    stack[sp++] = (size_t)(y + 1) * ms + x + 1;
                } else {
                    mrow[x] = 0;
                }
            }
        }
    }
    static const int ddx[8] = {-1, 0, 1, -1, 1, -1, 0, 1};
    static const int ddy[8] = {-1, -1, -1, 0, 0, 1, 1, 1};
    while (sp) {
        size_t p = stack[--sp];
        for (int k = 0; k < 8; ++k) {
            size_t q = p + (size_t)((long)ddy[k] * ms + ddx[k]);
            if (map[q] == 0) {
                map[q] = 2;
                stack[sp++] = q;
            }
        }
    }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x)
            dst[(size_t)y * w + x] = (map[(size_t)(y + 1) * ms + x + 1] == 2) ? 255 : 0;
    free(stack); free(map); free(mag); free(dx); free(dy);
}

/* cv::dilate / cv::erode with the default 3x3 rectangle, default border
 * (BORDER_CONSTANT with morphologyDefaultBorderValue: out-of-image pixels never
 * win the max / min). */
static void morph3_once(const uint8_t* src, int w, int h, int is_dilate, uint8_t* dst)
{
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            int v = is_dilate ? 0 : 255;
            for (int j = -1; j <= 1; ++j)
                for (int i = -1; i <= 1; ++i) {
                    int yy = y + j, xx = x + i;
                    if (yy < 0 || yy >= h || xx < 0 || xx >= w) continue;
                    int s = src[(size_t)yy * w + xx];
                    if (is_dilate ? (s > v) : (s < v)) v = s;
                }
            dst[(size_t)y * w + x] = (uint8_t)v;
        }
}

static void morph3(const uint8_t* src, int w, int h, int iters, int is_dilate, uint8_t* dst)
{
    size_t n = (size_t)w * h;
    uint8_t* a = (uint8_t*)malloc(n);
    uint8_t* b = (uint8_t*)malloc(n);
    memcpy(a, src, n);
    for (int i = 0; i < iters; ++i) {
        morph3_once(a, w, h, is_dilate, b);
        uint8_t* t = a; a = b; b = t;
    }
    memcpy(dst, a, n);
    free(a); free(b);
}

void mo_dilate3(const uint8_t* src, int w, int h, int iters, uint8_t* dst) { morph3(src, w, h, iters, 1, dst); }
void mo_erode3(const uint8_t* src, int w, int h, int iters, uint8_t* dst) { morph3(src, w, h, iters, 0, dst); }

void mo_edge_stage(const uint8_t* bgr, int w, int h, int stride, int canny_low,
                   uint8_t* canny_out, uint8_t* morph_out)
{
    size_t n = (size_t)w * h;
    uint8_t* a = (uint8_t*)malloc(n);
    uint8_t* b = (uint8_t*)malloc(n);
    mo_bgr2gray(bgr, w, h, stride, a);
    mo_blur3(a, w, h, b);
    mo_canny(b, w, h, canny_low, 3 * canny_low, a);
    if (canny_out) memcpy(canny_out, a, n);
    mo_dilate3(a, w, h, 2, b);
    mo_erode3(b, w, h, 1, morph_out);
    free(a); free(b);
}
