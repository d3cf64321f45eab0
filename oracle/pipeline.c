/*
 * pipeline.c -- oracle (TEST INFRASTRUCTURE ONLY): plain-C restatement of the mantis3
 * per-frame localization, quadDetection() in src/mantis3.cpp:68-135, and of the tf
 * (Bullet LinearMath) algebra it uses.  tf is a third-party dependency that is not in the
 * reference tree; the formulas below restate tf::Matrix3x3::setRotation / getRotation /
 * setRPY / getRPY and tf::Transform compose / inverse as published in ROS geometry
 * (tf/LinearMath).  The reference records no expected outputs for this path; the OpenCV
 * pieces are pinned against cv2 4.13, RPP against the compiled reference (see each file).
 *
 * Stated normalisations of reference behaviour that is undefined or non-deterministic:
 *   - cv::RNG rng(1) is a process-wide global that carries over between frames
 *     (Mantis3Params.h:87); here every frame starts from the seed (exact for the first frame).
 *   - rng.gaussian() calls inside one argument list (PoseAdjustment.h:15-16) are evaluated
 *     right-to-left, as GCC does on x86-64.
 *   - FLANN's randomised KD-tree with checks=32 (QuadDetection.h:130,143) is replaced by the
 *     exact radius search it approximates; unfilled result slots alias index 0 as in the
 *     reference's zero-filled std::vector<int>.
 *   - std::sort in getBestNHypotheses (HypothesisEvaluation.h:500) is unstable; a stable sort
 *     is used (ties keep input order).
 *   - pixel reads the reference does not bounds-check (HypothesisEvaluation.h:222,250) follow
 *     its linear addressing inside the image buffer and read 0 outside it.
 */
#include "pipeline.h"
#include "mantis_oracle.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

void mo_default_params(mo_params* p)
{
    p->canny_hysteresis = 50;
    p->polygon_epsilon = 10;
    p->search_radius_multiplier = 0.1;
    p->grid_spacing = 0.32;
    p->grid_size = 9;
    p->quad_stretch_factor = 1.3;
    p->max_quad_error = 0.5;
    p->max_angle_difference = 0.2;
    p->projection_bias = 1.1;
    p->color_projection_bias = 1.1;
    p->rot_sigma = 0.03;
    p->trans_sigma = 0.01;
    p->particles = 50;
    p->iterations = 10;
    p->rng_seed = 1;
}

/* ------------------------------------------------------------------ tf algebra */
static void mat_mul(const double* a, const double* b, double* r)
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            r[i * 3 + j] = a[i * 3] * b[j] + a[i * 3 + 1] * b[3 + j] + a[i * 3 + 2] * b[6 + j];
}
static void mat_vec(const double* a, const double* v, double* r)
{
    for (int i = 0; i < 3; ++i) r[i] = a[i * 3] * v[0] + a[i * 3 + 1] * v[1] + a[i * 3 + 2] * v[2];
}
/* tf::Transform::operator* : basis = a.basis*b.basis, origin = a(b.origin) */
static void tf_mul(const mo_tf* a, const mo_tf* b, mo_tf* r)
{
    mo_tf o;
    mat_mul(a->R, b->R, o.R);
    mat_vec(a->R, b->t, o.t);
    for (int k = 0; k < 3; ++k) o.t[k] += a->t[k];
    *r = o;
}
/* tf::Transform::inverse : (R^T, R^T * -t) */
static void tf_inv(const mo_tf* a, mo_tf* r)
{
    mo_tf o;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) o.R[i * 3 + j] = a->R[j * 3 + i];
    double nt[3] = {-a->t[0], -a->t[1], -a->t[2]};
    mat_vec(o.R, nt, o.t);
    *r = o;
}
/* tf::Matrix3x3::getRotation */
static void mat_to_quat(const double* m, double* q)
{
    double trace = m[0] + m[4] + m[8];
    double temp[4];
    if (trace > 0.0) {
        double s = sqrt(trace + 1.0);
        temp[3] = s * 0.5;
        s = 0.5 / s;
        temp[0] = (m[7] - m[5]) * s;
        temp[1] = (m[2] - m[6]) * s;
        temp[2] = (m[3] - m[1]) * s;
    } else {
        int i = m[0] < m[4] ? (m[4] < m[8] ? 2 : 1) : (m[0] < m[8] ? 2 : 0);
        int j = (i + 1) % 3, k = (i + 2) % 3;
        double s = sqrt(m[i * 3 + i] - m[j * 3 + j] - m[k * 3 + k] + 1.0);
        temp[i] = s * 0.5;
        s = 0.5 / s;
        temp[3] = (m[k * 3 + j] - m[j * 3 + k]) * s;
        temp[j] = (m[j * 3 + i] + m[i * 3 + j]) * s;
        temp[k] = (m[k * 3 + i] + m[i * 3 + k]) * s;
    }
    memcpy(q, temp, sizeof(temp));
}
/* tf::Matrix3x3::setRotation */
static void quat_to_mat(const double* q, double* m)
{
    double d = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
    double s = 2.0 / d;
    double xs = q[0] * s, ys = q[1] * s, zs = q[2] * s;
    double wx = q[3] * xs, wy = q[3] * ys, wz = q[3] * zs;
    double xx = q[0] * xs, xy = q[0] * ys, xz = q[0] * zs;
    double yy = q[1] * ys, yz = q[1] * zs, zz = q[2] * zs;
    m[0] = 1.0 - (yy + zz); m[1] = xy - wz;         m[2] = xz + wy;
    m[3] = xy + wz;         m[4] = 1.0 - (xx + zz); m[5] = yz - wx;
    m[6] = xz - wy;         m[7] = yz + wx;         m[8] = 1.0 - (xx + yy);
}
/* tf::Matrix3x3::setRPY == setEulerYPR(yaw, pitch, roll) */
void mo_tf_set_rpy(double roll, double pitch, double yaw, double* m)
{
    double ci = cos(roll), cj = cos(pitch), ch = cos(yaw);
    double si = sin(roll), sj = sin(pitch), sh = sin(yaw);
    double cc = ci * ch, cs = ci * sh, sc = si * ch, ss = si * sh;
    m[0] = cj * ch; m[1] = sj * sc - cs; m[2] = sj * cc + ss;
    m[3] = cj * sh; m[4] = sj * ss + cc; m[5] = sj * cs - sc;
    m[6] = -sj;     m[7] = cj * si;      m[8] = cj * ci;
}
/* tf::Matrix3x3::getRPY (solution 1) */
void mo_tf_get_rpy(const double* m, double* rpy)
{
    if (fabs(m[6]) >= 1) {
        double delta = atan2(m[7], m[8]);
        rpy[2] = 0;
        rpy[1] = (m[6] < 0) ? M_PI / 2.0 : -M_PI / 2.0;
        rpy[0] = delta;
    } else {
        double pitch = -asin(m[6]);
        rpy[1] = pitch;
        rpy[0] = atan2(m[7] / cos(pitch), m[8] / cos(pitch));
        rpy[2] = atan2(m[3] / cos(pitch), m[0] / cos(pitch));
    }
}

/* Hypothesis::setC2W / setW2C, Mantis3Types.h:68-80 */
void mo_hyp_set_c2w(mo_hyp* h, const mo_tf* c2w)
{
    h->c2w = *c2w;
    tf_inv(&h->c2w, &h->w2c);
    mat_to_quat(h->w2c.R, h->q);
}
void mo_hyp_set_w2c(mo_hyp* h, const mo_tf* w2c)
{
    h->w2c = *w2c;
    tf_inv(&h->w2c, &h->c2w);
    mat_to_quat(h->w2c.R, h->q);
}

static void rotz90(mo_tf* r)
{
    double q[4] = {0, 0, 1 / sqrt(2), 1 / sqrt(2)};
    quat_to_mat(q, r->R);
    r->t[0] = r->t[1] = r->t[2] = 0;
}

/* ------------------------------------------------------------------ quads */
typedef struct {
    int c[8];
    float cx, cy;
    double side;
    double test[8];
    int neighbor;
} mquad_t;

/* Quadrilateral ctor, QuadDetection.h:46-61 (float centre, float stretch, widened to double) */
static void quad_init(mquad_t* q, const int* c, double stretch)
{
    memcpy(q->c, c, sizeof(int) * 8);
    q->neighbor = 0;
    int dx = c[0] - c[2], dy = c[1] - c[3];
    q->side = sqrt((double)(dx * dx + dy * dy));
    float xs = 0, ys = 0;
    for (int k = 0; k < 4; ++k) { xs += (float)c[2 * k]; ys += (float)c[2 * k + 1]; }
    q->cx = xs / 4;
    q->cy = ys / 4;
    for (int k = 0; k < 4; ++k) {
        float ex = (float)c[2 * k] - q->cx, ey = (float)c[2 * k + 1] - q->cy;
        float mx = (float)((double)ex * stretch), my = (float)((double)ey * stretch);
        q->test[2 * k] = (double)(mx + q->cx);
        q->test[2 * k + 1] = (double)(my + q->cy);
    }
}

/* removeDuplicateQuads, QuadDetection.h:115-171 */
static int remove_duplicate_quads(mquad_t* q, int n, double mult)
{
    for (int e = 0; e < n; ++e) {
        if (q[e].neighbor) continue;
        float radius = (float)(mult * q[e].side);
        int idx[4] = {0, 0, 0, 0};
        float dst[4];
        int found = 0;
        for (int j = 0; j < n; ++j) {
            float ddx = q[e].cx - q[j].cx, ddy = q[e].cy - q[j].cy;
            float d = 0;
            d += ddx * ddx;
            d += ddy * ddy;
            if (!(d <= radius)) continue;
            /* insert by (dist, index), keep the 4 smallest */
            int pos = found < 4 ? found : 4;
            while (pos > 0 && (dst[pos - 1] > d)) --pos;
            if (pos >= 4) continue;
            int last = found < 4 ? found : 3;
            for (int k = last; k > pos; --k) { dst[k] = dst[k - 1]; idx[k] = idx[k - 1]; }
            dst[pos] = d;
            idx[pos] = j;
            if (found < 4) ++found;
        }
        for (int k = 1; k < 4; ++k) q[idx[k]].neighbor = 1; /* unfilled slots are 0 */
    }
    int m = 0;
    for (int e = 0; e < n; ++e)
        if (!q[e].neighbor) q[m++] = q[e];
    return m;
}

static mquad_t g_quads[MO_MAX_QUADS]; /* the oracle is single-threaded test infrastructure */

/* detectQuadrilaterals from the morphology output on, QuadDetection.h:216-228 */
int mo_detect_quads(const uint8_t* morph, int w, int h, const mo_params* prm, mo_trace* tr)
{
    mo_contours cs;
    mo_find_contours(morph, w, h, MO_RETR_CCOMP, &cs);
    tr->n_contours = cs.n;
    int nq = 0;
    int* buf = (int*)malloc(sizeof(int) * 2 * (size_t)(cs.npts + 4));
    for (int i = 0; i < cs.n; ++i) {
        int m = mo_approx_poly_dp(cs.pts + 2 * (size_t)cs.c[i].first, cs.c[i].count, (double)prm->polygon_epsilon, 1, buf);
        if (m == 4 && nq < MO_MAX_QUADS) quad_init(&g_quads[nq++], buf, prm->quad_stretch_factor);
    }
    free(buf);
    mo_free_contours(&cs);
    tr->n_quads_raw = nq;
    nq = remove_duplicate_quads(g_quads, nq, prm->search_radius_multiplier);
    tr->n_quads = nq;
    for (int i = 0; i < nq; ++i) {
        memcpy(tr->quad_corners[i], g_quads[i].c, sizeof(int) * 8);
        memcpy(tr->quad_test_px[i], g_quads[i].test, sizeof(double) * 8);
    }
    return nq;
}

/* ------------------------------------------------------------------ hypotheses */
/* generateCentralHypotheses, HypothesisGeneration.h:57-109.  Returns 1 and 4 hypotheses if
 * the quad passes. */
static int central_hypotheses(const double* norm_pts, const mo_params* prm, mo_hyp* out4, double* img_err_out)
{
    const double s = prm->grid_spacing / 2;
    /* gridSquarePossibilities, Mantis3Params.h:102-123 */
    const double poss[2][4][2] = {{{s, s}, {-s, s}, {-s, -s}, {s, -s}}, {{s, -s}, {-s, -s}, {-s, s}, {s, s}}};
    mo_hyp hyp;
    memset(&hyp, 0, sizeof(hyp));
    double error = DBL_MAX; /* the reference leaves this uninitialised if RPP never runs */
    for (int e = 0; e < 2; ++e) {
        double model[12], ip[12];
        for (int i = 0; i < 4; ++i) {
            model[i] = poss[e][i][0]; model[4 + i] = poss[e][i][1]; model[8 + i] = 0;
            ip[i] = norm_pts[2 * i]; ip[4 + i] = norm_pts[2 * i + 1]; ip[8 + i] = 1.0;
        }
        mo_tf c2w;
        int it;
        double oe, ie;
        if (!mo_rpp(model, ip, 4, c2w.R, c2w.t, &it, &oe, &ie)) {
            /* CoPlanarPoseEstimator.cpp:42-53 reads an empty Mat here (UB): reject the quad */
            *img_err_out = DBL_MAX;
            return 0;
        }
        error = ie;
        mo_hyp_set_c2w(&hyp, &c2w);
        if (hyp.w2c.t[2] >= 0) break;
    }
    *img_err_out = error;
    if (error > prm->max_quad_error) return 0;
    mo_tf rz;
    rotz90(&rz);
    out4[0] = hyp;
    for (int k = 1; k < 4; ++k) {
        mo_tf w2c;
        tf_mul(&rz, &out4[k - 1].w2c, &w2c);
        out4[k] = hyp;
        mo_hyp_set_w2c(&out4[k], &w2c);
    }
    return 1;
}

/* PoseClusterer::clusterByAngle + BFcluster + keepLargestCluster, PoseClusterer.cpp:33-116,
 * PoseClusterer.h:59-85.  Returns the member indices of the first-largest cluster. */
static int cluster_by_angle(const mo_hyp* hyps, int n, double rad, int* members)
{
    float (*ang)[3] = malloc(sizeof(float[3]) * (size_t)(n ? n : 1));
    char* taken = calloc((size_t)(n ? n : 1), 1);
    int* idx = malloc(sizeof(int) * (size_t)(n ? n : 1));
    for (int i = 0; i < n; ++i) {
        double m[9], rpy[3];
        quat_to_mat(hyps[i].q, m);
        mo_tf_get_rpy(m, rpy);
        for (int k = 0; k < 3; ++k) ang[i][k] = (float)rpy[k];
    }
    int best_n = 0, have = 0;
    for (int i = 0; i < n; ++i) {
        if (taken[i]) continue;
        int cnt = 0;
        for (int j = 0; j < n; ++j) {
            if (taken[j]) continue;
            double dx = ang[i][0] - ang[j][0], dy = ang[i][1] - ang[j][1], dz = ang[i][2] - ang[j][2];
            if (sqrt(dx * dx + dy * dy + dz * dz) <= rad * 1.5) idx[cnt++] = j;
        }
        float c[3] = {0, 0, 0};
        for (int k = 0; k < cnt; ++k) { c[0] += ang[idx[k]][0]; c[1] += ang[idx[k]][1]; c[2] += ang[idx[k]][2]; }
        for (int k = 0; k < 3; ++k) c[k] = (float)((double)c[k] / (double)cnt);
        cnt = 0;
        for (int j = 0; j < n; ++j) {
            if (taken[j]) continue;
            double dx = c[0] - ang[j][0], dy = c[1] - ang[j][1], dz = c[2] - ang[j][2];
            if (sqrt(dx * dx + dy * dy + dz * dz) <= rad) idx[cnt++] = j;
        }
        for (int k = 0; k < cnt; ++k) taken[idx[k]] = 1;
        if (!have || cnt > best_n) { /* keepLargestCluster: strictly larger replaces */
            if (!have || cnt > best_n) { best_n = cnt; memcpy(members, idx, sizeof(int) * (size_t)cnt); }
            have = 1;
        }
    }
    free(ang); free(taken); free(idx);
    return best_n;
}

/* ------------------------------------------------------------------ edge mask */
static void box_morph(const uint8_t* src, int w, int h, int k, int dilate, uint8_t* dst)
{
    /* (2k+1)^2 box max/min; pixels outside the image never win (== BORDER_REPLICATE here) */
    uint8_t* tmp = malloc((size_t)w * h);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            int v = dilate ? 0 : 255;
            int x0 = x - k < 0 ? 0 : x - k, x1 = x + k >= w ? w - 1 : x + k;
            for (int i = x0; i <= x1; ++i) {
                int s = src[(size_t)y * w + i];
                v = dilate ? (s > v ? s : v) : (s < v ? s : v);
            }
            tmp[(size_t)y * w + x] = (uint8_t)v;
        }
    for (int y = 0; y < h; ++y) {
        int y0 = y - k < 0 ? 0 : y - k, y1 = y + k >= h ? h - 1 : y + k;
        for (int x = 0; x < w; ++x) {
            int v = dilate ? 0 : 255;
            for (int j = y0; j <= y1; ++j) {
                int s = tmp[(size_t)j * w + x];
                v = dilate ? (s > v ? s : v) : (s < v ? s : v);
            }
            dst[(size_t)y * w + x] = (uint8_t)v;
        }
    }
    free(tmp);
}

/* cleanImageByEdge's mask from the Canny map, HypothesisEvaluation.h:336-372:
 * morphological gradient (3x3 cross), NOT, findContours(RETR_LIST)+drawContours(all, 255, 1)
 * (== every foreground pixel of the NOT image with a 4-neighbour that is background or
 * outside the image: pinned against cv2 by tests/test_oracle_pipeline.py), OR with Canny,
 * closes of 3,4,5 iterations, erode 3. */
void mo_clean_mask(const uint8_t* canny, int w, int h, uint8_t* mask)
{
    size_t n = (size_t)w * h;
    uint8_t* t = malloc(n);
    uint8_t* a = malloc(n);
#define AT(img, yy, xx, oob) (((yy) < 0 || (yy) >= h || (xx) < 0 || (xx) >= w) ? (oob) : (img)[(size_t)(yy) * w + (xx)])
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            int c = canny[(size_t)y * w + x];
            int dil = c | AT(canny, y - 1, x, 0) | AT(canny, y + 1, x, 0) | AT(canny, y, x - 1, 0) | AT(canny, y, x + 1, 0);
            int ero = c & AT(canny, y - 1, x, 255) & AT(canny, y + 1, x, 255) & AT(canny, y, x - 1, 255) & AT(canny, y, x + 1, 255);
            int grad = (dil && !ero) ? 255 : 0;
            t[(size_t)y * w + x] = (uint8_t)(255 - grad);
        }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            int v = t[(size_t)y * w + x];
            int interior = AT(t, y - 1, x, 0) && AT(t, y + 1, x, 0) && AT(t, y, x - 1, 0) && AT(t, y, x + 1, 0);
            int small = v && !interior;
            a[(size_t)y * w + x] = (uint8_t)((canny[(size_t)y * w + x] || small) ? 255 : 0);
        }
#undef AT
    for (int i = 0; i < 3; ++i) { /* MASK_CLOSE_ASCEND_ITER, MASK_INIT_CLOSE_ITER + i */
        box_morph(a, w, h, 3 + i, 1, t);
        box_morph(t, w, h, 3 + i, 0, a);
    }
    box_morph(a, w, h, 3, 0, mask); /* MASK_ERODE_ITER */
    free(t); free(a);
}

/* ------------------------------------------------------------------ scoring */
static int cv_round(double v) { return (int)lrint(v); }

/* img.at<Vec3b>(Point) with the reference's unchecked linear addressing.  The image the reference indexes is
 * the clone() of the message's image (src/mantis3.cpp:71) or the copyTo() output of cleanImageByEdge
 * (HypothesisEvaluation.h:384): both tightly packed, so pixel (x, y) is element y * w + x whatever x is
 * (x == w wraps into the next row); `stride` only says where the caller keeps row yy of that image. */
static int pixel_err(const uint8_t* bgr, const uint8_t* mask, int w, int h, int stride, int x, int y, const int* des)
{
    long pix = (long)y * w + x;
    int b = 0, g = 0, r = 0;
    if (pix >= 0 && pix < (long)w * h) {
        long yy = pix / w, xx = pix - yy * w;
        int keep = 1;
        if (mask) keep = mask[pix] != 0;
        if (keep) {
            const uint8_t* p = bgr + yy * stride + 3 * xx;
            b = p[0]; g = p[1]; r = p[2];
        }
    }
    int db = b - des[0], dg = g - des[1], dr = r - des[2];
    return db * db + dg * dg + dr * dr;
}

static int project(const mo_hyp* hyp, const double* p, const float* K, const double* D, int w, int h, double* px)
{
    double c[3];
    mat_vec(hyp->c2w.R, p, c);
    for (int k = 0; k < 3; ++k) c[k] += hyp->c2w.t[k];
    if (!(c[2] > 0)) return 0;
    double nrm[2] = {c[0] / c[2], c[1] / c[2]};
    mo_fisheye_distort(nrm, 1, K, D, px);
    return (px[0] < w && px[1] < h && px[0] >= 0 && px[1] > 0); /* inFrame, HypothesisEvaluation.h:388 */
}

/* evaluateHypothesis (fast path) on the edge-masked image, HypothesisEvaluation.h:71-158: colour error and number of
 * projected points, ADDED to *err / *n (computeImageError's `projected` is a running count over the cameras of a rig,
 * include/legacy/mantis2/Mantis2Error.h:32-53,55-99) */
static void accum_white(const mo_hyp* hyp, const uint8_t* bgr, const uint8_t* mask, int w, int h, int stride, const float* K,
                        const double* D, const mo_map* map, double* err, int* n)
{
    static const int WHITE[3] = {255, 255, 255};
    const double* sets[3] = {map->white, map->red, map->green};
    const int counts[3] = {map->n_white, map->n_red, map->n_green};
    for (int s = 0; s < 3; ++s)
        for (int i = 0; i < counts[s]; ++i) {
            double px[2];
            if (!project(hyp, sets[s] + 3 * i, K, D, w, h, px)) continue;
            (*n)++;
            *err += (double)pixel_err(bgr, mask, w, h, stride, cv_round(px[0]), cv_round(px[1]), WHITE);
        }
}

double mo_score_white(const mo_hyp* hyp, const uint8_t* bgr, const uint8_t* mask, int w, int h, int stride,
                      const float* K, const double* D, const mo_map* map, double bias)
{
    double err = 0;
    int n = 0;
    accum_white(hyp, bgr, mask, w, h, stride, K, D, map, &err, &n);
    if (n <= 0) return DBL_MAX;
    return err / ((double)n * bias);
}

/* evaluateHypothesisCOLOR (slow path, GREEN_ONLY) on the original image, :89-105,160-266 */
static void accum_green(const mo_hyp* hyp, const uint8_t* bgr, int w, int h, int stride, const float* K, const double* D,
                        const mo_map* map, double* err, int* n)
{
    static const int GREEN[3] = {50, 255, 85};
    for (int i = 0; i < map->n_green; ++i) {
        double px[2];
        if (!project(hyp, map->green + 3 * i, K, D, w, h, px)) continue;
        (*n)++;
        double e = 0;
        for (double x = -5.0; x < 5.0; x += 1)
            for (double y = -5.0; y < 5.0; y += 1)
                e += (double)pixel_err(bgr, NULL, w, h, stride, cv_round(px[0] + x), cv_round(px[1] + y), GREEN);
        e /= 100.0;
        *err += e;
    }
}

double mo_score_green(const mo_hyp* hyp, const uint8_t* bgr, int w, int h, int stride, const float* K,
                      const double* D, const mo_map* map, double bias)
{
    double err = 0;
    int n = 0;
    accum_green(hyp, bgr, w, h, stride, K, D, map, &err, &n);
    if (n <= 0) return DBL_MAX;
    return err / ((double)n * bias);
}

/* ------------------------------------------------------------------ selection / refinement */
/* getBestNHypotheses, HypothesisEvaluation.h:484-518: sort by error descending, keep the last
 * n.  Stable insertion sort (see header note). */
static int best_n(mo_hyp* hyps, int count, int n)
{
    if (count <= n) return count;
    for (int i = 1; i < count; ++i) {
        mo_hyp key = hyps[i];
        int j = i - 1;
        while (j >= 0 && hyps[j].error < key.error) { hyps[j + 1] = hyps[j]; --j; }
        hyps[j + 1] = key;
    }
    memmove(hyps, hyps + (count - n), sizeof(mo_hyp) * (size_t)n);
    return n;
}

/* One camera of the rig a hypothesis is scored on.  A hypothesis is the pose of the BASE in the world; camera k sees
 * it as W2C_k = W2B * b2c_k (Mantis2Error.h:67-68: new_hyp.setW2C(hyp.getW2C() * b2c); MantisRequest.h:14-26).
 * A single camera without extrinsics (mantis3's quadDetection) is the case n_cams == 1, has_b2c == 0. */
typedef struct {
    const uint8_t* bgr;
    const uint8_t* mask;
    const float* K;
    const double* D;
    mo_tf b2c;
    int has_b2c;
} eval_cam;

typedef struct {
    int n_cams;
    eval_cam cam[MO_MAX_RIG];
    int w, h, stride;
    const mo_map* map;
    const mo_params* prm;
    long scored;
} eval_ctx;

static void camera_hyp(const eval_cam* c, const mo_hyp* hyp, mo_hyp* out)
{
    if (!c->has_b2c) { *out = *hyp; return; }
    mo_tf w2c;
    tf_mul(&hyp->w2c, &c->b2c, &w2c);
    *out = *hyp;
    mo_hyp_set_w2c(out, &w2c);
}

/* joint score: summed colour error over the cameras / (summed projected points * bias) */
static double eval_white(eval_ctx* e, const mo_hyp* hyp)
{
    e->scored++;
    double err = 0;
    int n = 0;
    for (int k = 0; k < e->n_cams; ++k) {
        mo_hyp ch;
        camera_hyp(&e->cam[k], hyp, &ch);
        accum_white(&ch, e->cam[k].bgr, e->cam[k].mask, e->w, e->h, e->stride, e->cam[k].K, e->cam[k].D, e->map, &err, &n);
    }
    if (n <= 0) return DBL_MAX;
    return err / ((double)n * e->prm->projection_bias);
}

static double eval_green(eval_ctx* e, const mo_hyp* hyp)
{
    e->scored++;
    double err = 0;
    int n = 0;
    for (int k = 0; k < e->n_cams; ++k) {
        mo_hyp ch;
        camera_hyp(&e->cam[k], hyp, &ch);
        accum_green(&ch, e->cam[k].bgr, e->w, e->h, e->stride, e->cam[k].K, e->cam[k].D, e->map, &err, &n);
    }
    if (n <= 0) return DBL_MAX;
    return err / ((double)n * e->prm->color_projection_bias);
}

/* generateRandomHypothesis, PoseAdjustment.h:13-23 (argument lists evaluated right to left) */
static void random_hypothesis(const mo_hyp* hyp, const mo_params* prm, uint64_t* rng, mo_hyp* out)
{
    double yaw = mo_rng_gaussian(rng, prm->rot_sigma);
    double pitch = mo_rng_gaussian(rng, prm->rot_sigma);
    double roll = mo_rng_gaussian(rng, prm->rot_sigma);
    double z = mo_rng_gaussian(rng, prm->trans_sigma);
    double y = mo_rng_gaussian(rng, prm->trans_sigma);
    double x = mo_rng_gaussian(rng, prm->trans_sigma);
    mo_tf rnd, w2c;
    mo_tf_set_rpy(roll, pitch, yaw, rnd.R);
    rnd.t[0] = x; rnd.t[1] = y; rnd.t[2] = z;
    tf_mul(&hyp->w2c, &rnd, &w2c);
    *out = *hyp;
    mo_hyp_set_w2c(out, &w2c);
}

/* optimizeHypothesisWithParticleFilter, PoseAdjustment.h:29-60 */
static void optimize_hypothesis(eval_ctx* e, const mo_hyp* in, uint64_t* rng, mo_hyp* out)
{
    mo_hyp best = *in;
    best.error = eval_white(e, &best);
    for (int i = 0; i < e->prm->iterations; ++i) {
        mo_hyp sample = best;
        for (int j = 0; j < e->prm->particles; ++j) {
            mo_hyp tmp;
            random_hypothesis(&sample, e->prm, rng, &tmp);
            tmp.error = eval_white(e, &tmp);
            if (tmp.error < best.error) best = tmp;
        }
    }
    *out = best;
}

/* ------------------------------------------------------------------ the frame / the rig */
/* quadDetection() (src/mantis3.cpp:68-135) for n_cams cameras of one rig.  n_cams == 1 without extrinsics is the
 * reference's callback itself.  With extrinsics (cam_to_base: x y z qx qy qz qw per camera, the b2c of
 * MantisRequest.h:14-26) every camera contributes its quads' hypotheses, moved to the BASE frame
 * (W2B = W2C_k * b2c_k^-1), the rotation clustering runs on the pooled set, and every score from step (6) on is the
 * joint one of eval_white / eval_green: the multi-camera refinement the reference states as its intent
 * (MonteCarlo.cpp:233-241 "w2c2 ... TODO", Mantis2Error.h:55-99) and never finished.  tr's quad arrays hold the quads
 * of all cameras in camera order; the pose is the base's. */
static int localize_impl(int n_cams, const uint8_t* const* bgrs, int w, int h, int stride, const float* Ks, const double* Ds,
                         const double* cam_to_base, const mo_params* prm, const mo_map* map, mo_trace* tr, uint8_t* mask_out)
{
    memset(tr, 0, sizeof(*tr));
    if (n_cams < 1 || n_cams > MO_MAX_RIG) return -1;
    size_t n = (size_t)w * h;
    uint8_t* canny = malloc(n);
    uint8_t* morph = malloc(n);
    uint8_t* masks = malloc(n * (size_t)n_cams);
    mo_hyp* hyps = malloc(sizeof(mo_hyp) * (MO_MAX_HYPS + 4));
    mo_hyp* sel = malloc(sizeof(mo_hyp) * (MO_MAX_HYPS + 4));
    int* members = malloc(sizeof(int) * (MO_MAX_HYPS + 4));
    mo_trace* one = n_cams > 1 ? malloc(sizeof(mo_trace)) : tr;
    int status = MO_OK;
    eval_ctx e;
    memset(&e, 0, sizeof(e));
    e.n_cams = n_cams; e.w = w; e.h = h; e.stride = stride; e.map = map; e.prm = prm;

    int nq_total = 0, nh = 0;
    for (int k = 0; k < n_cams; ++k) {
        const float* K = Ks + 9 * k;
        const double* D = Ds + 4 * k;
        eval_cam* c = &e.cam[k];
        c->bgr = bgrs[k]; c->K = K; c->D = D; c->mask = masks + n * (size_t)k;
        c->has_b2c = cam_to_base != NULL;
        mo_tf b2c_inv;
        if (c->has_b2c) {
            const double* q = cam_to_base + 7 * k;
            quat_to_mat(q + 3, c->b2c.R);
            memcpy(c->b2c.t, q, sizeof(double) * 3);
            tf_inv(&c->b2c, &b2c_inv);
        }
        /* (1) detectQuadrilaterals, src/mantis3.cpp:79 */
        mo_edge_stage(bgrs[k], w, h, stride, prm->canny_hysteresis, canny, morph);
        /* (5) cleanImageByEdge, :102 (needs this camera's Canny map; used from step (6) on) */
        mo_clean_mask(canny, w, h, masks + n * (size_t)k);
        if (n_cams > 1) memset(one, 0, sizeof(*one));
        int nq = mo_detect_quads(morph, w, h, prm, one);
        /* (2) undistortAndNormalizeQuadTestPoints + (3) generateHypotheses, :87 */
        for (int i = 0; i < nq; ++i) {
            mo_fisheye_undistort(one->quad_test_px[i], 4, K, D, one->quad_test_norm[i]);
            mo_hyp four[4];
            int pass = central_hypotheses(one->quad_test_norm[i], prm, four, &one->quad_img_err[i]);
            one->quad_pass[i] = pass;
            if (!pass) continue;
            if (c->has_b2c)
                for (int j = 0; j < 4; ++j) {   /* camera pose -> base pose */
                    mo_tf w2b;
                    tf_mul(&four[j].w2c, &b2c_inv, &w2b);
                    mo_hyp_set_w2c(&four[j], &w2b);
                }
            if (nh + 4 <= MO_MAX_HYPS) { memcpy(hyps + nh, four, sizeof(four)); nh += 4; }
        }
        if (n_cams > 1) {   /* append this camera's quads to the rig's trace */
            tr->n_contours += one->n_contours;
            tr->n_quads_raw += one->n_quads_raw;
            for (int i = 0; i < nq && nq_total + i < MO_MAX_QUADS; ++i) {
                memcpy(tr->quad_corners[nq_total + i], one->quad_corners[i], sizeof(one->quad_corners[i]));
                memcpy(tr->quad_test_px[nq_total + i], one->quad_test_px[i], sizeof(one->quad_test_px[i]));
                memcpy(tr->quad_test_norm[nq_total + i], one->quad_test_norm[i], sizeof(one->quad_test_norm[i]));
                tr->quad_pass[nq_total + i] = one->quad_pass[i];
                tr->quad_img_err[nq_total + i] = one->quad_img_err[i];
            }
        }
        nq_total += nq;
    }
    tr->n_quads = nq_total;
    if (mask_out) memcpy(mask_out, masks, n);
    if (nq_total == 0) { status = MO_NO_QUADS; goto done; }
    tr->n_hyps_generated = nh;

    /* (4) cluster by angle, keep the largest cluster, :90-92 */
    int nc = cluster_by_angle(hyps, nh, prm->max_angle_difference, members);
    for (int k = 0; k < nc; ++k) sel[k] = hyps[members[k]];
    tr->n_hyps_cluster = nc;
    if (nc == 0) { status = MO_NO_HYPOTHESES; goto done; }

    /* (6) evaluateHypotheses (7) best 1, :107-110 */
    for (int k = 0; k < nc; ++k) sel[k].error = eval_white(&e, &sel[k]);
    nc = best_n(sel, nc, 1);
    tr->first_best = sel[nc - 1];

    /* (8) optimizeHypothesisWithParticleFilter, :113 */
    uint64_t rng = prm->rng_seed ? prm->rng_seed : 0xffffffffULL;
    mo_hyp optim;
    optimize_hypothesis(&e, &sel[nc - 1], &rng, &optim);
    tr->optimized = optim;

    /* (9) computeAllShiftedHypothesesFAST, HypothesisGeneration.h:125-140 (10) evaluate */
    int ns = 0;
    {
        double gs = prm->grid_spacing, half = ((double)prm->grid_size / 2.0) * gs;
        for (double x = -half + gs / 2.0; x < half; x += gs)
            for (double y = -half + gs / 2.0; y < half; y += gs) {
                mo_tf w2c = optim.w2c;
                w2c.t[0] += x; w2c.t[1] += y; w2c.t[2] += 0;
                memset(&hyps[ns], 0, sizeof(mo_hyp));
                mo_hyp_set_w2c(&hyps[ns], &w2c);
                hyps[ns].error = eval_white(&e, &hyps[ns]);
                if (ns < 81) tr->shift_err[ns] = hyps[ns].error;
                ns++;
            }
    }
    /* (11) best 20, :123 */
    ns = best_n(hyps, ns, 20);

    /* (12) determineBestYaw on the original image, HypothesisEvaluation.h:521-581 */
    {
        mo_tf rz;
        rotz90(&rz);
        mo_hyp* rots[4];
        for (int r = 0; r < 4; ++r) rots[r] = malloc(sizeof(mo_hyp) * (size_t)ns);
        memcpy(rots[0], hyps, sizeof(mo_hyp) * (size_t)ns);
        for (int r = 1; r < 4; ++r)
            for (int j = 0; j < ns; ++j) {
                mo_tf w2c;
                tf_mul(&rz, &rots[r - 1][j].w2c, &w2c);
                rots[r][j] = rots[r - 1][j];
                mo_hyp_set_w2c(&rots[r][j], &w2c);
            }
        double best_error = DBL_MAX;
        int best = -1;
        for (int r = 0; r < 4; ++r) {
            double total = 0;
            int success = 0;
            for (int j = 0; j < ns; ++j) {
                rots[r][j].error = eval_green(&e, &rots[r][j]);
                if (fabs(rots[r][j].error - DBL_MAX) > 0.001) { success++; total += rots[r][j].error; }
            }
            tr->yaw_err[r] = success ? total / (double)success : DBL_MAX;
            if (tr->yaw_err[r] < best_error) { best_error = tr->yaw_err[r]; best = r; }
        }
        double min1 = DBL_MAX, min2 = DBL_MAX;
        for (int r = 0; r < 4; ++r) {
            double diff = tr->yaw_err[r] - best_error;
            if (diff < min1) { min2 = min1; min1 = diff; }
            else if (diff < min2) min2 = diff;
        }
        tr->min_yaw_diff = min2;
        tr->best_rot = best;
        if (best < 0) status = MO_YAW_FAILED; /* the reference calls back() on an empty vector here */
        else tr->final_hyp = rots[best][ns - 1];
        for (int r = 0; r < 4; ++r) free(rots[r]);
    }
    if (status == MO_OK) {
        memcpy(tr->pos, tr->final_hyp.w2c.t, sizeof(double) * 3);
        memcpy(tr->quat_xyzw, tr->final_hyp.q, sizeof(double) * 4);
        tr->error = tr->final_hyp.error;
    }
    tr->n_scored = e.scored;
done:
    tr->status = status;
    if (n_cams > 1) free(one);
    free(canny); free(morph); free(masks); free(hyps); free(sel); free(members);
    return status;
}

int mo_localize(const uint8_t* bgr, int w, int h, int stride, const float* K, const double* D,
                const mo_params* prm, const mo_map* map, mo_trace* tr, uint8_t* mask_out)
{
    return localize_impl(1, &bgr, w, h, stride, K, D, NULL, prm, map, tr, mask_out);
}

int mo_localize_rig(const uint8_t* const* bgr, int n_cams, int w, int h, int stride, const float* K, const double* D,
                    const double* cam_to_base, const mo_params* prm, const mo_map* map, mo_trace* tr)
{
    return localize_impl(n_cams, bgr, w, h, stride, K, D, cam_to_base, prm, map, tr, NULL);
}
