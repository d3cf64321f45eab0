/*
 * mantis_oracle.h -- CPU oracle for the mantis localization hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a plain-C restatement of the arithmetic the
 * reference (k-sheridan/mantis, mantis3 generation) performs per frame, including
 * the OpenCV 4.13 fixed-point kernels it calls.  It exists to check the CUDA
 * library (libmantis_cuda.so) and to be timed as the CPU baseline.  Nothing in the
 * product path may include, link or call it.
 *
 * Parity pin: every function here is checked against cv2 4.13.0 (the OpenCV the
 * reference's calls resolve to in this environment) on test/grid1.jpg,
 * test/lena.jpg and seeded noise by tests/test_oracle_*.py, with golden digests
 * committed under tests/golden/.  The reference itself has no assertions or
 * recorded outputs for these stages.
 *
 * Each function cites the reference call site it restates (file:line relative to
 * the reference tree).
 */
#ifndef MANTIS_ORACLE_H
#define MANTIS_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- image stage: include/mantis3/QuadDetection.h:209-214 ---- */
void mo_bgr2gray(const uint8_t* bgr, int w, int h, int stride, uint8_t* gray);
void mo_blur3(const uint8_t* src, int w, int h, uint8_t* dst);
void mo_sobel3(const uint8_t* src, int w, int h, int16_t* dx, int16_t* dy);
void mo_canny(const uint8_t* src, int w, int h, int low, int high, uint8_t* dst);
void mo_dilate3(const uint8_t* src, int w, int h, int iters, uint8_t* dst);
void mo_erode3(const uint8_t* src, int w, int h, int iters, uint8_t* dst);
/* whole chain of detectQuadrilaterals up to the contour input */
void mo_edge_stage(const uint8_t* bgr, int w, int h, int stride, int canny_low,
                   uint8_t* canny_out /* may be NULL */, uint8_t* morph_out);

/* ---- contours: QuadDetection.h:216-226, HypothesisEvaluation.h:343 ---- */
enum { MO_RETR_LIST = 1, MO_RETR_CCOMP = 2 };
typedef struct mo_contour {
    int first, count; /* slice of mo_contours.pts (points, not ints) */
    int is_hole;
    int parent;       /* discovery index of the enclosing border, -1 = frame */
    int discovery;    /* raster discovery index of this border */
} mo_contour;
typedef struct mo_contours {
    int n, npts;
    mo_contour* c;
    int* pts; /* x,y pairs */
} mo_contours;
int mo_find_contours(const uint8_t* src, int w, int h, int mode, mo_contours* out);
void mo_free_contours(mo_contours* c);
/* cv::approxPolyDP on an int contour; dst holds up to n points; returns the new count */
int mo_approx_poly_dp(const int* src_xy, int n, double eps, int closed, int* dst_xy);

/* ---- camera model + RNG: QuadDetection.h:294, Mantis3Types.h:125-136, PoseAdjustment.h:13-23 ---- */
void mo_fisheye_undistort(const double* src_xy, int n, const float* K, const double* D, double* dst_xy);
void mo_fisheye_distort(const double* src_xy, int n, const float* K, const double* D, double* dst_xy);
void mo_ziggurat_tables(unsigned* kn, float* wn, float* fn);
float mo_rng_randn(uint64_t* state);
double mo_rng_gaussian(uint64_t* state, double sigma);

/* ---- planar pose: CoPlanarPoseEstimator.cpp:16-58 -> RPP::Rpp (RPP.cpp:13-64) ---- */
#define MO_RPP_MAX_POINTS 16
int mo_rpp(const double* model3xn, const double* iprts3xn, int n, double* R9, double* t3, int* iters,
           double* obj_err, double* img_err);

/* ---- Gauss-Newton / PnP in its mantis2 meaning: include/legacy/mantis2/PoseEstimator.h:16-25,91-118 ---- */
void mo_rodrigues_exp(const double* w, double* R);
void mo_rodrigues_log(const double* R, double* w);
int mo_pnp_iterative(const float* obj, const float* img, int m, const float* K, const double* guess6, int flags,
                     int max_iters, double* out9);

/* ---- legacy particle weight, fisheye::undistortImage, Markov yaw filter (oracle/legacy.c) ---- */
void mo_markov_update_weights(double* yaw360, double stddev);
void mo_markov_init(double* p360, double yaw_rad);
void mo_markov_convolve(double* p360, double dtheta, double dt);
void mo_markov_sense(double* p360, double yaw_rad);
void mo_markov_update_errors(const double* p360, const double* yaw_rad, double* err, int n);
double mo_markov_get_yaw(const double* p360, int* bin);
double mo_camera_error(const uint8_t* img, int w, int h, int stride, const float* K, const double* w2c12, const double* map,
                       int n_white, int n_red, int n_green);
void mo_fisheye_undistort_maps(const float* K, const double* D, int w, int h, short* map1, unsigned short* map2);
void mo_remap_linear_bgr(const uint8_t* src, int w, int h, int stride, const short* map1, const unsigned short* map2,
                         uint8_t* dst, int dst_stride);

/* ---- synthetic frame source (stands in for GridRenderer.cpp:211-353) ---- */
void mo_render_grid(const double* R_w2c, const double* t_w2c, const float* K, const double* D, int w, int h,
                    double line_width, int floor_level, uint8_t* bgr);

#ifdef __cplusplus
}
#endif
#endif
