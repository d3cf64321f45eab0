/*
 * mantis_cuda.h -- C ABI of libmantis_cuda.so: the B200 (sm_100a) implementation of
 * mantis's per-frame localization hot path.
 *
 * The reference (k-sheridan/mantis) has no plugin / FFI layer: its stages are
 * header-only free functions compiled into the ROS node.  The seam is therefore
 * placed directly under its two entry points
 *   - topic callback  quadDetection(ImageConstPtr, CameraInfoConstPtr)  src/mantis3.cpp:68
 *   - service handler runMantis(Request&, Response&)                    src/mantis_server.cpp:23
 * mantis_localize_batch() takes what those receive (images + K + D) and returns
 * what they emit (pose, error/weight, counts; srv/mantisService.srv:1-12).  The
 * stage-level entry points expose each reference function on the path so that
 * tests and benchmarks can check and time it in isolation.
 *
 * Conventions
 *   - every function returns MANTIS_OK (0) or a negative mantis_status; nothing
 *     throws or exits across the ABI (the reference exit(1)s inside RPP.cpp:452).
 *   - `mem` says where the caller's image / result buffers live: MANTIS_MEM_HOST
 *     (the library copies straight from / to the caller's buffers on its streams --
 *     pin them with cudaHostAlloc / cudaHostRegister if the copy is to overlap the
 *     kernels; pageable memory works, synchronously) or MANTIS_MEM_DEVICE (pointers
 *     are device pointers on the context's device).
 *   - a context is bound to one device and one stream; one context per host
 *     thread.  No allocation happens on the hot calls once the context has seen
 *     the largest batch / frame size (scratch grows monotonically).
 *   - all calls are synchronous with respect to the host unless stated.
 */
#ifndef MANTIS_CUDA_H
#define MANTIS_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mantis_ctx mantis_ctx;

typedef enum mantis_status {
    MANTIS_OK = 0,
    MANTIS_ERR_INVALID = -1,     /* bad argument */
    MANTIS_ERR_CUDA = -2,        /* CUDA runtime failure; see mantis_last_error() */
    MANTIS_ERR_NO_DEVICE = -3,   /* no CUDA device: there is no CPU fallback */
    MANTIS_ERR_UNSUPPORTED = -4, /* frame too large for the on-chip stages */
    MANTIS_ERR_NOMEM = -5
} mantis_status;

enum { MANTIS_MEM_HOST = 0, MANTIS_MEM_DEVICE = 1 };

/* Tunables of the reference: rosparams of getParameters()
 * (include/mantis3/Mantis3Params.h:154-194) and its compile-time defines (:12-68). */
typedef struct mantis_config {
    int32_t canny_hysteresis;   /* ~canny_hysteresis, default 50; Canny(T, 3T)   :162 */
    int32_t polygon_epsilon;    /* ~polygon_epsilon, default 10                  :164 */
    double search_radius_multiplier; /* ~neighborhood_search_radius_multiplier .1 :166 */
    double grid_spacing;        /* ~grid_spacing, default 0.32                   :167 */
    int32_t grid_size;          /* GRID_SIZE 9                                   :23  */
    double quad_stretch_factor; /* QUAD_STRECH_FACTOR 1.3                        :20  */
    double max_quad_error;      /* MAX_QUAD_ERROR 0.5                            :26  */
    double max_angle_difference;/* MAX_ANGLE_DIFFERENCE 0.2                      :28  */
    double projection_bias;     /* PROJECTION_BIAS 1.1                           :37  */
    double color_projection_bias; /* COLOR_PROJECTION_BIAS 1.1                   :55  */
    double rot_sigma;           /* ROT_SIGMA 0.03                                :62  */
    double trans_sigma;         /* TRANS_SIGMA 0.01                              :63  */
    int32_t particles;          /* optimizeHypothesisWithParticleFilter default 50  PoseAdjustment.h:29 */
    int32_t iterations;         /* ... default 10 */
    uint64_t rng_seed;          /* cv::RNG rng(1)                                :87  */
} mantis_config;

/* Fills *cfg with the reference's defaults. */
void mantis_default_config(mantis_config* cfg);

int mantis_create(const mantis_config* cfg, int device, mantis_ctx** out);
int mantis_destroy(mantis_ctx* ctx);
const char* mantis_strerror(int status);
/* Text of the last CUDA error seen by this context (empty string if none). */
const char* mantis_last_error(const mantis_ctx* ctx);
/* Number of kernels this context has launched since creation (bench.py's gpu_launches). */
uint64_t mantis_launch_count(const mantis_ctx* ctx);
/* cudaStream_t of the context, as an opaque pointer (for event timing on the launching stream). */
void* mantis_stream(const mantis_ctx* ctx);

/*
 * Per-stage device timing of mantis_localize_batch (CUDA events on the context stream).
 * Stages: 0 input staging (H2D), 1 candidate kernel (gray/blur/Sobel/NMS), 2 edge-link kernel
 * (hysteresis, morphology, mask), 3 quad kernel, 4 pose kernel, 5 search kernel, 6 result D2H.
 */
#define MANTIS_N_STAGES 7
int mantis_set_profiling(mantis_ctx* ctx, int enabled);
/* Copies the accumulated milliseconds per stage into ms[MANTIS_N_STAGES] and resets them. */
int mantis_get_stage_ms(mantis_ctx* ctx, double* ms);

/*
 * Image stage of detectQuadrilaterals(): cvtColor(BGR2GRAY) -> GaussianBlur(3x3, s=3)
 * -> Canny(T, 3T, 3) -> dilate x2 -> erode x1   (include/mantis3/QuadDetection.h:209-214;
 * the first three are repeated verbatim by cleanImageByEdge, HypothesisEvaluation.h:323-327).
 *
 *   bgr          n frames of 8-bit BGR, HWC; row r of frame f starts at
 *                bgr + f*frame_pitch_bytes + r*stride_bytes
 *   canny_out    optional (NULL to skip): n*height*width bytes, {0,255}, the cv::Canny result
 *   morph_out    optional: n*height*width bytes, {0,255}, after dilate x2 / erode x1
 */
int mantis_edges_batch(mantis_ctx* ctx, const uint8_t* bgr, int n, int width, int height,
                       int stride_bytes, size_t frame_pitch_bytes,
                       uint8_t* canny_out, uint8_t* morph_out, int mem);

/*
 * Edge mask of cleanImageByEdge() (include/mantis3/HypothesisEvaluation.h:319-386): the image the
 * WHITE scoring reads is `bgr where mask else 0`; the library keeps the mask as one bit per pixel
 * and never materialises the masked copy.  mask_out: n*height*width bytes {0,255}.
 */
int mantis_mask_batch(mantis_ctx* ctx, const uint8_t* bgr, int n, int width, int height,
                      int stride_bytes, size_t frame_pitch_bytes, uint8_t* mask_out, int mem);

/*
 * Grid-cell (quad) extraction, the rest of detectQuadrilaterals():
 * findContours(RETR_CCOMP, CHAIN_APPROX_SIMPLE) -> approxPolyDP(eps, closed) -> keep 4-gons ->
 * Quadrilateral ctor -> removeDuplicateQuads   (include/mantis3/QuadDetection.h:13-66,115-171,216-228).
 * Quads come out in the reference's order (OpenCV's RETR_CCOMP contour order).
 */
#define MANTIS_MAX_QUADS 256
enum { /* mantis_frame_quads.flags: capacity limits of the on-chip stage hit for this frame */
    MANTIS_FLAG_BORDER_OVERFLOW = 1, /* more than 1024 contours */
    MANTIS_FLAG_POINT_OVERFLOW = 2,  /* more than 65536 contour vertices */
    MANTIS_FLAG_QUAD_OVERFLOW = 4    /* more than MANTIS_MAX_QUADS 4-gons */
};
typedef struct mantis_frame_quads {
    int32_t n_contours;   /* contours found (f->contours.size()) */
    int32_t n_quads_raw;  /* 4-gons before removeDuplicateQuads */
    int32_t n_quads;      /* f->quads.size() */
    int32_t flags;
    int32_t corners[MANTIS_MAX_QUADS][8]; /* Quadrilateral::contour, x,y of 4 vertices */
    double test_px[MANTIS_MAX_QUADS][8];  /* Quadrilateral::test_points (pixels, before undistortion) */
} mantis_frame_quads;

int mantis_quads_batch(mantis_ctx* ctx, const uint8_t* bgr, int n, int width, int height,
                       int stride_bytes, size_t frame_pitch_bytes, mantis_frame_quads* out, int mem);

/*
 * Line voting on the stage's edge map (north-star stage "Hough / line-voting accumulator, line
 * clustering, grid-intersection extraction").  The reference itself has no Hough transform -- its
 * grid primitives are the cell quads of mantis_quads_batch (include/mantis3/QuadDetection.h:216-226) --
 * so this entry point is not on the localization path.  Voting and peak picking follow
 * cv::HoughLines(edges, rho = 1, theta = pi/180, threshold) of the OpenCV the reference links
 * (accumulator (180+2) x (2(W+H)+1+2), 4-neighbour local maxima above threshold, strongest first,
 * ties by accumulator index).  Clustering and intersections are defined by this library
 * (oracle/hough.c states them): greedy (rho, theta) clustering around the strongest lines with
 * vote-weighted means, then the crossings of the two dominant pencils that fall inside the image.
 */
#define MANTIS_HOUGH_MAX_LINES 1024
#define MANTIS_HOUGH_MAX_CLUSTERS 256
#define MANTIS_HOUGH_MAX_POINTS 1024
enum { MANTIS_HOUGH_FLAG_PEAK_OVERFLOW = 1 };  /* more local maxima than the peak buffer holds (4096) */
typedef struct mantis_hough_line { float rho, theta; int32_t votes; } mantis_hough_line;
typedef struct mantis_hough_cluster { float rho, theta; int32_t votes, members; } mantis_hough_cluster;
typedef struct mantis_frame_lines {
    int32_t n_peaks;     /* local maxima above threshold (what cv::HoughLines would return) */
    int32_t n_lines;     /* min(n_peaks, MANTIS_HOUGH_MAX_LINES) */
    int32_t n_clusters, n_points, flags, pad_;
    mantis_hough_line lines[MANTIS_HOUGH_MAX_LINES];
    mantis_hough_cluster clusters[MANTIS_HOUGH_MAX_CLUSTERS];
    float points[MANTIS_HOUGH_MAX_POINTS][2];
} mantis_frame_lines;
typedef struct mantis_hough_params {
    int32_t threshold;     /* accumulator threshold (votes > threshold) */
    int32_t use_morph;     /* 0: vote on the Canny map, 1: on the dilate x2 / erode x1 map */
    double rho_tol;        /* clustering tolerances */
    double theta_tol;
    double pencil_tol;     /* angular half-width of a pencil */
} mantis_hough_params;
int mantis_hough_batch(mantis_ctx* ctx, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                       size_t frame_pitch_bytes, const mantis_hough_params* prm, mantis_frame_lines* out, int mem);

/*
 * Gauss-Newton / Levenberg-Marquardt PnP, batched: the mantis2 generation's pose step
 *   computeHypothesis -> cv::solvePnP(obj, img, K, noArray, rvec, tvec, useExtrinsicGuess, SOLVEPNP_ITERATIVE)
 *                                                   include/legacy/mantis2/PoseEstimator.h:91-118
 *   error = computeHypothesisQuadReprojectionError  :16-25   (sqrt of the summed squared pixel residuals)
 * One pose per frame from m correspondences (pinhole K, no distortion, as the legacy call passes noArray).
 *   obj_xyz   m*3 floats per frame (or one shared set when obj_shared != 0)
 *   img_xy    m*2 floats per frame (pixels)
 *   K         9 floats per frame, row-major (or one shared K when k_shared != 0)
 *   guess_rt  optional: rvec(3) + tvec(3) doubles per frame = useExtrinsicGuess; NULL = planar (z = 0)
 *             initialisation from a homography, as OpenCV does for planar objects
 * The result is the minimiser of the reprojection error in the basin of the initial pose (OpenCV's LM
 * converges to the same point; its iteration path is not reproduced).
 */
enum { MANTIS_PNP_BOTH_PLANAR = 1 };  /* params.flags: also refine the mirrored planar candidate, keep the better (cf. RPP.cpp:36-52) */
enum { MANTIS_PNP_OK = 0, MANTIS_PNP_DEGENERATE = 1 };
typedef struct mantis_pnp_params {
    int32_t flags;
    int32_t max_iters;  /* LM iterations per candidate; <= 0 means 50 */
} mantis_pnp_params;
typedef struct mantis_pnp_result {
    double rvec[3];       /* Rodrigues vector, cv::Rodrigues' branch conventions */
    double tvec[3];
    double reproj_error;  /* computeHypothesisQuadReprojectionError */
    int32_t passes;       /* evaluations of the normal equations */
    int32_t status;
} mantis_pnp_result;
int mantis_pnp_batch(mantis_ctx* ctx, const float* obj_xyz, const float* img_xy, int n, int m, int obj_shared,
                     const float* K, int k_shared, const double* guess_rt, const mantis_pnp_params* prm,
                     mantis_pnp_result* out, int mem);

/*
 * Planar pose from point correspondences, batched: RPP::Rpp(model_3D, iprts, rotation, translation, iterations,
 * obj_err, img_err) (include/mantis3/RobustPlanarPose/RPP.h:54-55, RPP.cpp:13-64) as called per quad by
 * CoPlanarPoseEstimator::estimatePose (CoPlanarPoseEstimator.cpp:16-58).
 *   model_xyz   m x (x, y, z) doubles per problem (one shared set when model_shared != 0)
 *   image_xy    m x (x, y) doubles per problem: NORMALISED image points (the reference passes z = 1 rows)
 *   m           3 .. 16 points.  Planar 4-point problems (z == 0: every quad of the localization path) run through
 *               the very kernels of mantis_localize_batch; other shapes through the general 3x3-SVD form.
 * img_err carries the reference's quirk (every point compared with image point 0, RPP.cpp:197-198): it is the
 * quantity generateCentralHypotheses gates on (HypothesisGeneration.h:80).
 */
enum { MANTIS_RPP_OK = 0, MANTIS_RPP_FAILED = 1 };  /* FAILED: RPP::Rpp returned false (no second-pose candidate) */
typedef struct mantis_rpp_result {
    double R[9];            /* rotation (row-major) and translation: p_camera = R p_model + t */
    double t[3];
    double obj_err, img_err;
    int32_t iters_first;    /* AbsKernel calls of the first ObjPose run (RPP.cpp:66-208); -1 = not recorded */
    int32_t iters_best;     /* ... of the refinement run that won */
    int32_t n_candidates;   /* second-pose candidates refined (Get2ndPose_Exact, RPP.cpp:693-753) */
    int32_t status;
} mantis_rpp_result;
int mantis_rpp_batch(mantis_ctx* ctx, const double* model_xyz, int model_shared, const double* image_xy, int n, int m,
                     mantis_rpp_result* out, int mem);

/*
 * Map-correspondence score of given poses on given frames, batched: evaluateHypothesisWithImageWHITE on the image
 * cleanImageByEdge produces (include/mantis3/HypothesisEvaluation.h:107-158, 319-386) -- steps (5) + (6) of
 * quadDetection() for caller-supplied hypotheses, through the scoring loop of mantis_localize_batch.
 *   K, D      9 floats + 4 doubles per frame (as in mantis_frame)
 *   w2c       n x n_hyps x 12 doubles: R row-major + t of each hypothesis' W2C (what Hypothesis::setW2C receives)
 *   scores    n x n_hyps doubles: sum of squared colour distances to WHITE / (projected points x PROJECTION_BIAS),
 *             DBL_MAX when no map point projects into the frame
 */
int mantis_score_batch(mantis_ctx* ctx, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                       size_t frame_pitch_bytes, const float* K, const double* D, const double* w2c, int n_hyps,
                       double* scores, int mem);

/*
 * Whole-frame localization: everything quadDetection() does between receiving the image and
 * calling publishPose() (src/mantis3.cpp:68-135).
 */
typedef struct mantis_frame {     /* one sensor_msgs/Image + CameraInfo of mantisService.srv:1-2 */
    const uint8_t* bgr;           /* 8-bit BGR, HWC (host or device, see `mem`) */
    int32_t width, height, stride_bytes;
    float K[9];                   /* CameraInfo.K as the CV_32F 3x3 of get3x3FromVector, QuadDetection.h:189-201; K[1] (skew) must be 0 */
    double D[4];                  /* CameraInfo.D, fisheye coefficients (src/mantis3.cpp:74) */
    double cam_to_base[7];        /* camera->base_link transform (x y z qx qy qz qw), read by mantis_localize_rig_batch; all-zero quaternion = identity */
    uint64_t stamp_ns;
    int32_t camera_id;
} mantis_frame;

enum { /* mantis_result.status */
    MANTIS_FRAME_OK = 0,
    MANTIS_FRAME_NO_QUADS = 1,       /* src/mantis3.cpp:83-84 early return */
    MANTIS_FRAME_NO_HYPOTHESES = 2,  /* src/mantis3.cpp:94-97 early return */
    MANTIS_FRAME_YAW_FAILED = 3,     /* determineBestYaw found no finite score (the reference calls back() on an empty vector) */
    MANTIS_FRAME_OVERFLOW = 4        /* a capacity limit of the on-chip stages was hit (see MANTIS_FLAG_*) */
};

typedef struct mantis_result {    /* what publishPose() emits (PosePub.h:38-56) + the srv response fields */
    double pos[3];                /* hyp.getPosition(): origin of W2C */
    double quat_xyzw[4];          /* hyp.getQuaternion() */
    double error;                 /* hyp.error (variance = error/600, PosePub.h:47); srv `weight` */
    double min_yaw_diff;          /* publishPose gates on this > MINIMUM_YAW_DIFFERENCE (4000), PosePub.h:16 */
    int32_t status;
    int32_t n_quads;
    int32_t n_hyps;               /* hypotheses kept by the rotation clustering */
    int32_t n_particles;          /* hypotheses scored for this frame; srv `num_particles` */
} mantis_result;

typedef struct mantis_frame_trace { /* intermediate values, for parity tests */
    int32_t n_contours, n_quads_raw, n_quads, flags;
    int32_t n_hyps_generated, n_hyps_cluster, best_rot, pad_;
    double first_best_error;      /* getBestNHypotheses(1) */
    double optimized_error;       /* after optimizeHypothesisWithParticleFilter */
    double optimized_w2c[12];     /* R row-major (9) + t (3) */
    double shift_err[81];
    double yaw_err[4];
} mantis_frame_trace;

/* World map (params/map.yaml): xyz triples, white first, then red, then green. */
int mantis_set_map(mantis_ctx* ctx, const double* xyz, int n_white, int n_red, int n_green);

/* frames[i].bgr lives where `mem` says; all frames of one call must share width/height/stride.
 * out[n] (and the optional trace[n]) are host arrays. */
int mantis_localize_batch(mantis_ctx* ctx, const mantis_frame* frames, int n, mantis_result* out,
                          mantis_frame_trace* trace, int mem);

/*
 * Asynchronous form of mantis_localize_batch: returns as soon as the whole batch (all its chunks) is queued on the
 * context's streams, so one host thread can keep several devices busy, or prepare the next batch meanwhile.
 * frames' pixels, out and trace must stay valid -- and the context idle otherwise -- until mantis_wait(ctx) has
 * returned MANTIS_OK; only then are out / trace filled.  With pageable (non-pinned) out / trace the runtime makes
 * the result copies synchronous and the call degrades gracefully towards the blocking one.  Errors detected while
 * queueing are returned at once (the streams are drained first); mantis_wait reports CUDA errors raised later.
 * Not available while per-stage profiling is on.  (The reference's callback is synchronous: src/mantis3.cpp:68-135
 * runs on the ros::spin() thread, :155.)
 */
int mantis_localize_batch_async(mantis_ctx* ctx, const mantis_frame* frames, int n, mantis_result* out,
                                mantis_frame_trace* trace, int mem);
int mantis_wait(mantis_ctx* ctx);

/*
 * Rigs: joint multi-camera localization (BASELINE.json configs[2]).  frames holds n_rigs x cams_per_rig records, the
 * cameras of a rig next to each other, each with its cam_to_base -- the per-camera base->camera transform of the
 * reference's requests (tf::StampedTransform b2c, include/legacy/mantis/MantisRequest.h:14-26).  Every camera runs
 * steps (1)-(3) of quadDetection() on its own image; its hypotheses are moved to the BASE frame
 * (W2B = W2C_k * b2c_k^-1) and pooled; from the clustering on, a hypothesis is a base pose and every score is the
 * joint one the reference states as its intent: camera k sees W2C_k = W2B * b2c_k (new_hyp.setW2C(hyp.getW2C() * b2c),
 * include/legacy/mantis2/Mantis2Error.h:67-68), colour errors and projected-point counts run on over the cameras
 * (computeImageError's `projected`, :55-99; MonteCarlo::computeWeight leaves the second camera as a TODO,
 * include/legacy/mantis/MonteCarlo.cpp:233-241).  Refinement, cell shifts and the yaw decision act on the base pose.
 * out / trace: one record per RIG (pose of the base in the world; n_quads etc. summed over the cameras).  A rig of one
 * camera at the base origin gives mantis_localize_batch's record.  The pooled hypotheses of a rig share the capacity
 * cams_per_rig x 1024.
 */
#define MANTIS_MAX_RIG_CAMERAS 8
int mantis_localize_rig_batch(mantis_ctx* ctx, const mantis_frame* frames, int n_rigs, int cams_per_rig, mantis_result* out,
                              mantis_frame_trace* trace, int mem);

/*
 * The 2016 generation's particle weight, batched over particles x frames:
 *   MonteCarlo::computeCameraError / project2d / computeWeight   include/legacy/mantis/MonteCarlo.cpp:169-241
 * Pinhole projection of the context's map (white, red, green) by every world->camera transform onto an
 * UNDISTORTED image; colour SSE against WHITE (255,255,255), RED (113,48,32), GREEN (50,100,37)
 * (MonteCarlo.h:53-55); fewer than 10 projected points -> 1e17 / count, as written.
 *   K         9 floats per frame (CameraInfo.K narrowed to CV_32F, MonteCarlo.cpp:148-161)
 *   w2c       per frame, per particle: R row-major (9) + t (3) doubles -- w2c1 of computeWeight (:234-236)
 *   weights   n * n_particles doubles
 */
int mantis_particle_weights_batch(mantis_ctx* ctx, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                                  size_t frame_pitch_bytes, const float* K, const double* w2c, int n_particles,
                                  double* weights, int mem);

/* cv::fisheye::undistortImage(img, out, K, D, Knew = K) as parseRequest calls it (MonteCarlo.cpp:284): one
 * camera (upper-triangular K, 4 fisheye coefficients) for all n frames; out is n tightly packed BGR frames. */
int mantis_undistort_batch(mantis_ctx* ctx, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                           size_t frame_pitch_bytes, const float* K, const double* D, uint8_t* out, int mem);

/*
 * Temporal yaw filter, one 360-bin plane per camera stream: MarkovModel of
 * include/mantis3/Markov.cpp.  planes: n_streams x 360 doubles, updated in place.
 *   MANTIS_MARKOV_INIT      a = yaw (rad) of the first hypothesis' W2C: MarkovModel(Hypothesis)      :14-26
 *   MANTIS_MARKOV_CONVOLVE  a = dTheta (rad), b = dt (s): convolve()                                    :227-258
 *   MANTIS_MARKOV_SENSE     a = yaw (rad) of the accepted hypothesis: senseFusion(Hypothesis)           :164-205
 * mantis_markov_update_errors scales hypothesis errors by 1/p[yaw bin] of their stream: updateHypothesis() :207-220
 */
enum { MANTIS_MARKOV_INIT = 0, MANTIS_MARKOV_CONVOLVE = 1, MANTIS_MARKOV_SENSE = 2 };
int mantis_markov_batch(mantis_ctx* ctx, double* planes, int n_streams, int op, const double* a, const double* b, int mem);
int mantis_markov_update_errors(mantis_ctx* ctx, const double* planes, int n_streams, const int* stream_of,
                                const double* yaw, double* err, int n_hyps, int mem);

/*
 * Synthetic frame source on the device: the floor grid of params/map.yaml ray-cast through the
 * fisheye model for n camera poses (stands in for GridRenderer::renderGrid,
 * include/mantis3/GridRenderer.cpp:292-353, whose OpenGL path cannot run here).
 *   poses_w2c   host array, per frame 12 doubles: R (row-major, camera->world) then t (camera origin)
 *   d_bgr_out   DEVICE buffer receiving n frames of 8-bit BGR
 *   noise_sigma > 0 adds N(0, sigma) per channel from a counter-based hash of (seed, frame, pixel)
 */
int mantis_render_grid_batch(mantis_ctx* ctx, const double* poses_w2c, int n, int width, int height,
                             const float* K, const double* D, double line_width, int floor_level,
                             float noise_sigma, uint32_t seed, uint8_t* d_bgr_out, int stride_bytes,
                             size_t frame_pitch_bytes);

/*
 * BASELINE.json configs[3]'s frame degradation on the device ("1920x1080 synthetic grid frames with noise/blur"):
 * separable Gaussian blur (7 taps, sigma in pixels, BORDER_REFLECT_101; sigma 0 = none), then additive
 * N(0, noise_sigma) per channel (counter-based hash of seed, frame, pixel), rounded and clipped to 8 bits.
 * Both buffers are DEVICE memory and must not alias.  A frame source, like mantis_render_grid_batch; the
 * reference has no counterpart (its frames come from a camera).
 */
int mantis_degrade_batch(mantis_ctx* ctx, const uint8_t* d_bgr_in, uint8_t* d_bgr_out, int n, int width, int height,
                         int stride_in, size_t pitch_in, int stride_out, size_t pitch_out, float blur_sigma,
                         float noise_sigma, uint32_t seed);

/*
 * The image stage of mantis_edges_batch() for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  This entry point has NO counterpart in the
 * reference.  Same kernels, same results as
 * mantis_edges_batch(..., MANTIS_MEM_DEVICE), but the intermediate weak / strong candidate bit planes live in
 * a fixed file-scope __device__ array instead of context scratch, so the call performs no cudaMalloc /
 * cudaFree and calls no allocating library routine.
 *   - d_bgr, d_canny_out, d_morph_out are DEVICE pointers on the context's device (an output may be NULL,
 *     not both); layouts as in mantis_edges_batch.
 *   - limit: 2 * round_up_256(n * height * ceil(width / 32) * 4) <= MANTIS_NOALLOC_SCRATCH_BYTES (8 MiB), i.e.
 *     up to 109 frames of 640x480, 36 of 1280x720, 16 of 1920x1080 per call.  A larger batch is rejected
 *     with MANTIS_ERR_INVALID and a message in mantis_last_error(); split it, or use mantis_edges_batch.
 *   - every GPU that loads the library holds the 8 MiB array until the process exits.
 *   - all calls on one device SHARE the array, and it keeps the previous call's contents: the part a call
 *     uses is cleared on the context's stream first.  Two calls that overlap on the GPU (two contexts or
 *     streams on one device) race on it -- serialise them.
 * Synchronous with respect to the host, like the other entry points.
 */
#define MANTIS_NOALLOC_SCRATCH_BYTES ((size_t)8 << 20)
int mantis_edges_noalloc(mantis_ctx* ctx, const uint8_t* d_bgr, int n, int width, int height, int stride_bytes,
                         size_t frame_pitch_bytes, uint8_t* d_canny_out, uint8_t* d_morph_out);

#ifdef __cplusplus
}
#endif
#endif /* MANTIS_CUDA_H */
