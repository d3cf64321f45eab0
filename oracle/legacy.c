/*
 * legacy.c -- oracle (TEST INFRASTRUCTURE ONLY, see mantis_oracle.h) for the SURVEY 8(f) rows 3 and 4:
 *
 *  (3) the 2016 generation's particle weight
 *        MonteCarlo::project2d / computeCameraError      include/legacy/mantis/MonteCarlo.cpp:169-226
 *        colours WHITE / RED / GREEN                      include/legacy/mantis/MonteCarlo.h:53-55
 *      and the image it is evaluated on, cv::fisheye::undistortImage(img, K, D, K)   MonteCarlo.cpp:284
 *      (third-party: OpenCV fisheye.cpp initUndistortRectifyMap with CV_16SC2 maps + cv::remap INTER_LINEAR,
 *      BORDER_CONSTANT; restated here and checked against cv2 4.13 in tests/test_legacy.py),
 *  (4) the temporal yaw filter MarkovModel                include/mantis3/Markov.cpp:9-274
 */
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "mantis_oracle.h"

#define DEG 360
#define MO_PI 3.14159265358979323846

/* ---- Markov.cpp:28-31 ---- */
static double markov_weight(double x, double mu, double stddev, double y)
{
    return (y / (stddev * sqrt(2 * MO_PI))) * exp(-((x - mu) * (x - mu) / (2 * stddev * stddev)));
}

static void markov_normalize(const double* in, double* out)
{
    double sum = 0.0;
    for (int i = 0; i < DEG; ++i) sum += in[i];
    for (int i = 0; i < DEG; ++i) out[i] = in[i] / sum;
}

/* MarkovModel::updateWeights, Markov.cpp:76-124 (loop bounds and wrap-around exactly as written) */
void mo_markov_update_weights(double* yaw, double stddev)
{
    double aux[DEG];
    for (int i = 0; i < DEG; ++i) {
        double max = 0.0;
        int diff = (DEG / 2) + i;
        if (i <= DEG / 2) {
            for (int j = 0; j < diff; ++j) max += markov_weight((double)i, (double)j, stddev, yaw[j]);
            for (int j = diff; j < DEG; ++j) max += markov_weight((double)i, (double)(-(DEG - j)), stddev, yaw[j]);
        } else {
            for (int j = i; j < diff; ++j) max += markov_weight((double)i, (double)j, stddev, yaw[j % DEG]);
            for (int j = (diff % DEG); j < i; ++j) max += markov_weight((double)i, (double)j, stddev, yaw[j]);
        }
        aux[i] = max;
    }
    markov_normalize(aux, yaw);
}

static int yaw_bin(double yaw_rad)
{
    double y = yaw_rad * 180.0 / MO_PI;
    if (y < 0) y += DEG;
    int b = (int)y;
    return b < 0 ? 0 : (b >= DEG ? DEG - 1 : b);  /* the reference indexes unchecked; y = 360.0 cannot occur for yaw in (-pi, pi] */
}

/* MarkovModel(Hypothesis), Markov.cpp:14-26: delta at the yaw bin of W2C, smoothed with stddev 3 */
void mo_markov_init(double* p, double yaw_rad)
{
    for (int i = 0; i < DEG; ++i) p[i] = 0.0;
    p[yaw_bin(yaw_rad)] = 1;
    mo_markov_update_weights(p, 3);
}

/* MarkovModel::convolve, Markov.cpp:227-258.  `(int)dTheta*180/M_PI` casts dTheta (radians) first. */
void mo_markov_convolve(double* p, double dtheta, double dt)
{
    double aux[DEG];
    int disp = (int)((int)dtheta * 180 / MO_PI);
    if (dtheta > 0) {
        for (int i = disp; i < DEG + disp; ++i) aux[i % DEG] = p[i - disp];
    } else {
        disp *= -1;
        for (int i = 0; i < DEG; ++i) aux[i] = p[(disp + i) % DEG];
    }
    memcpy(p, aux, sizeof(aux));
    mo_markov_update_weights(p, (1.0 / 3.0) * dt * 11.5 / 30.0);
}

/* MarkovModel::senseFusion(Hypothesis) -> senseFusion(markovPlane), Markov.cpp:164-205.  The reference's
 * zeroing loop (`for(int i=0; sense.size(); ++i)`) never terminates; the evident intent (all zeros) is used. */
void mo_markov_sense(double* p, double yaw_rad)
{
    double sense[DEG];
    for (int i = 0; i < DEG; ++i) sense[i] = 0.0;
    sense[yaw_bin(yaw_rad)] = 1;
    mo_markov_update_weights(sense, 3.5);
    const double big = sqrt(DBL_MAX);
    for (int i = 0; i < DEG; ++i) { p[i] = p[i] * big; sense[i] = sense[i] * big; }
    for (int i = 0; i < DEG; ++i) p[i] *= sense[i];
    double tmp[DEG];
    markov_normalize(p, tmp);
    memcpy(p, tmp, sizeof(tmp));
}

/* MarkovModel::updateHypothesis, Markov.cpp:207-220 */
void mo_markov_update_errors(const double* p, const double* yaw_rad, double* err, int n)
{
    for (int i = 0; i < n; ++i) err[i] = err[i] * 1 / (p[yaw_bin(yaw_rad[i])]);
}

/* MarkovModel::getYaw, Markov.cpp:265-274 (returns the peak PROBABILITY scaled by pi/180, as written) */
double mo_markov_get_yaw(const double* p, int* bin)
{
    int max = 0;
    for (int i = 0; i < DEG; ++i)
        if (p[i] > p[max]) max = i;
    if (bin) *bin = max;
    return p[max] * MO_PI / 180;
}

/* ---- computeCameraError, MonteCarlo.cpp:183-226 ---- */
static int cv_round(double v) { return (int)lrint(v); }

double mo_camera_error(const uint8_t* img, int w, int h, int stride, const float* K, const double* w2c12, const double* map,
                       int n_white, int n_red, int n_green)
{
    static const int COL[3][3] = {{255, 255, 255}, {113, 48, 32}, {50, 100, 37}}; /* WHITE, RED, GREEN (MonteCarlo.h:53-55) */
    const int counts[3] = {n_white, n_red, n_green};
    int projected = 0, idx = 0;
    double error = 0;
    for (int set = 0; set < 3; ++set)
        for (int k = 0; k < counts[set]; ++k, ++idx) {
            const double* p = map + 3 * idx;
            double c[3];
            for (int i = 0; i < 3; ++i)
                c[i] = w2c12[i * 3] * p[0] + w2c12[i * 3 + 1] * p[1] + w2c12[i * 3 + 2] * p[2] + w2c12[9 + i];
            /* project2d: float K entries times double coordinates, result narrowed to Point2f */
            const float x = (float)(((double)K[0] * c[0]) / c[2] + (double)K[2]);
            const float y = (float)(((double)K[4] * c[1]) / c[2] + (double)K[5]);
            if (x > 0 && x < (float)w && y > 0 && y < (float)h) {
                /* img.at<Vec3b>(Point2f): Point2f -> Point rounds (cvRound); unchecked linear addressing */
                const long long pix = (long long)cv_round((double)y) * w + cv_round((double)x);
                int b = 0, g = 0, r = 0;
                if (pix >= 0 && pix < (long long)w * h) {
                    const uint8_t* q = img + (size_t)(pix / w) * stride + 3 * (pix % w);
                    b = q[0]; g = q[1]; r = q[2];
                }
                error += (double)((b - COL[set][0]) * (b - COL[set][0])) + (double)((g - COL[set][1]) * (g - COL[set][1])) +
                         (double)((r - COL[set][2]) * (r - COL[set][2]));
                projected++;
            }
        }
    if (projected < 10) error = 1e17;
    return error / (double)projected;
}

/* ---- cv::fisheye::undistortImage(src, dst, K, D, Knew = K) ---- */
/* iR = (Knew * I)^-1 for an upper-triangular camera matrix, closed form */
static void inv_upper3(const double* K, double* iR)
{
    const double fx = K[0], s = K[1], cx = K[2], fy = K[4], cy = K[5];
    iR[0] = 1.0 / fx; iR[1] = -s / (fx * fy); iR[2] = (s * cy - cx * fy) / (fx * fy);
    iR[3] = 0;        iR[4] = 1.0 / fy;       iR[5] = -cy / fy;
    iR[6] = 0;        iR[7] = 0;              iR[8] = 1;
}

/* map1: w*h pairs of short (x, y); map2: w*h ushort (5-bit fractions, y*32 + x) */
void mo_fisheye_undistort_maps(const float* Kf, const double* D, int w, int h, short* map1, unsigned short* map2)
{
    double K[9], iR[9];
    for (int i = 0; i < 9; ++i) K[i] = Kf[i];
    inv_upper3(K, iR);
    const double f0 = K[0], f1 = K[4], c0 = K[2], c1 = K[5];
    for (int i = 0; i < h; ++i) {
        double _x = i * iR[1] + iR[2], _y = i * iR[4] + iR[5], _w = i * iR[7] + iR[8];
        for (int j = 0; j < w; ++j) {
            double u, v;
            if (_w <= 0) {
                u = (_x > 0) ? -INFINITY : INFINITY;
                v = (_y > 0) ? -INFINITY : INFINITY;
            } else {
                double x = _x / _w, y = _y / _w;
                double r = sqrt(x * x + y * y);
                double theta = atan(r);
                double theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
                double theta_d = theta * (1 + D[0] * theta2 + D[1] * theta4 + D[2] * theta6 + D[3] * theta8);
                double scale = (r == 0) ? 1.0 : theta_d / r;
                u = f0 * x * scale + c0;
                v = f1 * y * scale + c1;
            }
            double us = u * 32, vs = v * 32;
            int iu = us >= 2147483647.0 ? 2147483647 : (us <= -2147483648.0 ? (-2147483647 - 1) : (int)lrint(us));
            int iv = vs >= 2147483647.0 ? 2147483647 : (vs <= -2147483648.0 ? (-2147483647 - 1) : (int)lrint(vs));
            map1[((size_t)i * w + j) * 2] = (short)(iu >> 5);
            map1[((size_t)i * w + j) * 2 + 1] = (short)(iv >> 5);
            map2[(size_t)i * w + j] = (unsigned short)((iv & 31) * 32 + (iu & 31));
            _x += iR[0]; _y += iR[3]; _w += iR[6];
        }
    }
}

/* cv::remap(src, dst, map1 (CV_16SC2), map2 (CV_16UC1), INTER_LINEAR, BORDER_CONSTANT 0) on 8UC3.
 * Fixed-point bilinear weights of 5-bit fractions: w = a*b*32 (sum 32768), result (sum + 16384) >> 15.
 * (The table entry of fraction (0,0) is {32767,0,0,1} in OpenCV; it yields the same pixel.) */
void mo_remap_linear_bgr(const uint8_t* src, int w, int h, int stride, const short* map1, const unsigned short* map2,
                         uint8_t* dst, int dst_stride)
{
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            const int sx = map1[((size_t)i * w + j) * 2], sy = map1[((size_t)i * w + j) * 2 + 1];
            const int fr = map2[(size_t)i * w + j] & 1023, fx = fr & 31, fy = fr >> 5;
            const int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
            uint8_t* d = dst + (size_t)i * dst_stride + 3 * j;
            for (int ch = 0; ch < 3; ++ch) {
                int p[4];
                for (int t = 0; t < 4; ++t) {
                    const int xx = sx + (t & 1), yy = sy + (t >> 1);
                    p[t] = (xx >= 0 && xx < w && yy >= 0 && yy < h) ? src[(size_t)yy * stride + 3 * xx + ch] : 0;
                }
                d[ch] = (uint8_t)((p[0] * w00 + p[1] * w01 + p[2] * w10 + p[3] * w11 + (1 << 14)) >> 15);
            }
        }
}
