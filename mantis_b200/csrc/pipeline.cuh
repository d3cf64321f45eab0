// pipeline.cuh -- device-side records shared by the per-frame kernels (quads, pose, search).
#pragma once
#include <stdint.h>

#include "../../include/mantis_cuda.h"
#define MANTIS_MAX_BORDERS 1024   // border starts (contours) per frame
#define MANTIS_MAX_PTS 65536      // CHAIN_APPROX_SIMPLE vertices per frame
#define MANTIS_MAX_HYPS (4 * MANTIS_MAX_QUADS)

enum {
    FRAME_FLAG_BORDER_OVERFLOW = MANTIS_FLAG_BORDER_OVERFLOW,
    FRAME_FLAG_POINT_OVERFLOW = MANTIS_FLAG_POINT_OVERFLOW,
    FRAME_FLAG_QUAD_OVERFLOW = MANTIS_FLAG_QUAD_OVERFLOW,
};

// Output of the quad kernel, input of the pose kernel (same record as the C ABI exposes).
typedef mantis_frame_quads FrameQuads;

// tf::Transform
struct Tf {
    double R[9];
    double t[3];
};

// Hypothesis (Mantis3Types.h:26-161): both directions + quaternion of w2c + error
struct Hyp {
    Tf c2w, w2c;
    double q[4];
    double error;
};

// Per-frame by-products of the pose kernel (also surfaced through mantis_frame_trace).
struct FramePose {
    int n_hyps_generated, n_sel;
    int quad_pass[MANTIS_MAX_QUADS];
    double quad_img_err[MANTIS_MAX_QUADS];
};
