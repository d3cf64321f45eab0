/*
 * pipeline.h -- oracle (TEST INFRASTRUCTURE ONLY, see mantis_oracle.h): data types of the
 * per-frame localization restatement, quadDetection() in src/mantis3.cpp:68-135.
 */
#ifndef MANTIS_ORACLE_PIPELINE_H
#define MANTIS_ORACLE_PIPELINE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MO_MAX_QUADS 1024
#define MO_MAX_HYPS (4 * MO_MAX_QUADS)
#define MO_MAX_RIG 8 /* cameras per rig */

/* tf::Transform: basis (row-major) + origin */
typedef struct mo_tf { double R[9]; double t[3]; } mo_tf;

/* Hypothesis, include/mantis3/Mantis3Types.h:26-161 */
typedef struct mo_hyp {
    mo_tf c2w, w2c;
    double q[4]; /* x y z w of w2c's rotation */
    double error;
} mo_hyp;

/* rosparams + defines, include/mantis3/Mantis3Params.h:12-68,154-194 */
typedef struct mo_params {
    int canny_hysteresis, polygon_epsilon;
    double search_radius_multiplier, grid_spacing;
    int grid_size;
    double quad_stretch_factor, max_quad_error, max_angle_difference, projection_bias, color_projection_bias;
    double rot_sigma, trans_sigma;
    int particles, iterations;
    uint64_t rng_seed;
} mo_params;
void mo_default_params(mo_params* p);

typedef struct mo_map { /* xyz triples; params/map.yaml */
    int n_white, n_red, n_green;
    const double* white;
    const double* red;
    const double* green;
} mo_map;

enum { MO_OK = 0, MO_NO_QUADS = 1, MO_NO_HYPOTHESES = 2, MO_YAW_FAILED = 3 };

typedef struct mo_trace {
    int status;
    int n_contours, n_quads_raw, n_quads;
    int quad_corners[MO_MAX_QUADS][8];  /* x,y of the 4 approxPolyDP vertices, after de-duplication */
    double quad_test_px[MO_MAX_QUADS][8];
    double quad_test_norm[MO_MAX_QUADS][8];
    int n_hyps_generated, n_hyps_cluster;
    int quad_pass[MO_MAX_QUADS];        /* generateCentralHypotheses accepted this quad */
    double quad_img_err[MO_MAX_QUADS];
    mo_hyp first_best;   /* getBestNHypotheses(1) */
    mo_hyp optimized;    /* after optimizeHypothesisWithParticleFilter */
    double shift_err[81];
    double yaw_err[4];
    int best_rot;
    double min_yaw_diff;
    mo_hyp final_hyp;    /* hyps.back() handed to publishPose */
    double pos[3], quat_xyzw[4], error;
    long n_scored;       /* hypotheses scored (the srv's num_particles) */
} mo_trace;

/* Whole frame.  K is the CV_32F 3x3 the reference builds (QuadDetection.h:189-201); D 4 doubles.
 * mask_out (optional, w*h) receives cleanImageByEdge's mask {0,255}. */
int mo_localize(const uint8_t* bgr, int w, int h, int stride, const float* K, const double* D,
                const mo_params* prm, const mo_map* map, mo_trace* tr, uint8_t* mask_out);

/* A rig of n_cams cameras (one image, K (9), D (4) and camera->base transform (x y z qx qy qz qw) each) localized
 * jointly: see localize_impl in pipeline.c.  The pose in tr is the BASE's. */
int mo_localize_rig(const uint8_t* const* bgr, int n_cams, int w, int h, int stride, const float* K, const double* D,
                    const double* cam_to_base, const mo_params* prm, const mo_map* map, mo_trace* tr);

/* pieces exposed for unit checks */
void mo_clean_mask(const uint8_t* canny, int w, int h, uint8_t* mask);
int mo_detect_quads(const uint8_t* morph, int w, int h, const mo_params* prm, mo_trace* tr);
double mo_score_white(const mo_hyp* hyp, const uint8_t* bgr, const uint8_t* mask, int w, int h, int stride,
                      const float* K, const double* D, const mo_map* map, double bias);
double mo_score_green(const mo_hyp* hyp, const uint8_t* bgr, int w, int h, int stride, const float* K,
                      const double* D, const mo_map* map, double bias);
void mo_hyp_set_c2w(mo_hyp* h, const mo_tf* c2w);
void mo_hyp_set_w2c(mo_hyp* h, const mo_tf* w2c);
void mo_tf_set_rpy(double roll, double pitch, double yaw, double* R9);
void mo_tf_get_rpy(const double* R9, double* rpy);

#ifdef __cplusplus
}
#endif
#endif
