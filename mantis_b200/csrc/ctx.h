// ctx.h -- context object behind the C ABI (include/mantis_cuda.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <vector>
#include "../../include/mantis_cuda.h"
#include "pipeline.cuh"

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct mantis_ctx {
    mantis_config cfg;
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    uint64_t launches = 0;
    char err[256] = {0};

    // optional per-stage timing (CUDA events on the context stream)
    void* dbg_clocks = nullptr;  // device array of 16 long long, set by mantis_debug_clocks
    bool profiling = false;
    cudaEvent_t ev[MANTIS_N_STAGES + 1] = {};
    double stage_ms[MANTIS_N_STAGES] = {};

    // scratch (grows monotonically)
    DevBuf weak, strong;      // bit planes from the candidate kernel
    DevBuf canny_bits, morph_bits;  // bit-packed results of the per-frame kernel
    DevBuf q_pts, q_dst, q_stack, q_owner, q_pad, q_tmp;   // quad kernel scratch
    DevBuf frame_quads;                      // FrameQuads[n]
    DevBuf quad_hyp, pose_offs;
    DevBuf cams, hyps, sel, pose_info, results, traces, mask_bits;  // localize scratch (per chunk)
    DevBuf render_poses;
    DevBuf leg_a, leg_b, leg_c;              // scratch of the legacy-generation entry points (legacy.cu)
    int n_red = 0;
    DevBuf pnp_stage;                        // host-caller staging of mantis_pnp_batch
    DevBuf hough_trig, hough_count, hough_peaks, hough_lines;  // line-voting stage (hough.cu)
    DevBuf d_map, d_rand, d_shifts;          // world map, perturbation table, cell shifts
    DevBuf d_shift_pts, d_shift_idx;         // the shift hypotheses through the map's lattice (search.cu: score_points)
    int n_shift_pts = 0;                     // 0: the map is no lattice for these shifts, every shift is projected on its own
    std::vector<double> host_shifts;         // [n_shift][2]
    int n_map = 0, n_green = 0, n_shift = 0;
    // two-lane execution of large device-resident batches (localize.cu): child contexts with their own
    // streams and scratch, so that the latency-bound kernels of one half overlap the other half's
    static constexpr int MAX_LANES = 4;
    mantis_ctx* lane[MAX_LANES] = {};
    bool is_lane = false;
    std::vector<double> host_map;            // what mantis_set_map received (replayed into the lanes)
    int map_planar = 0;                      // every map point has z == 0 (the reference's floor grid)
    DevBuf d_atan;                           // (i pi/256, tan(i pi/256)) table of the fast fp64 atan (search.cu)
    Tf rotz;                                 // +90 degrees about z
    DevBuf rig_tf;                           // rigs: per frame b2c and its inverse (localize.cu)
    DevBuf search_ticket;                    // k_search: frame ticket counter of the persistent CTAs
    DevBuf stage_in, stage_in2, stage_out0, stage_out1;  // device staging for MANTIS_MEM_HOST callers
    cudaStream_t copy_stream = nullptr;       // H2D of the next chunk while the current one computes
    cudaEvent_t ev_copied[2] = {}, ev_free[2] = {};
};

#define MCHECK(ctx, call)                                                                  \
    do {                                                                                   \
        cudaError_t e__ = (call);                                                          \
        if (e__ != cudaSuccess) {                                                          \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d %s: %s", __FILE__, __LINE__, #call, \
                     cudaGetErrorString(e__));                                             \
            return MANTIS_ERR_CUDA;                                                        \
        }                                                                                  \
    } while (0)

// stage timing helpers: mark(i) records the event that ends stage i-1 / starts stage i
inline void prof_mark(mantis_ctx* c, int i)
{
    if (c->profiling) cudaEventRecord(c->ev[i], c->stream);
}
inline void prof_collect(mantis_ctx* c, int first, int last)  // after a stream sync
{
    if (!c->profiling) return;
    for (int i = first; i < last; ++i) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, c->ev[i], c->ev[i + 1]) == cudaSuccess) c->stage_ms[i] += ms;
    }
}

inline int ctx_reserve(mantis_ctx* c, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return MANTIS_OK;
    if (b.p) { MCHECK(c, cudaFree(b.p)); b.p = nullptr; b.cap = 0; }
    MCHECK(c, cudaMalloc(&b.p, bytes));
    b.cap = bytes;
    return MANTIS_OK;
}

inline int ctx_reserve_pinned(mantis_ctx* c, void** p, size_t* cap, size_t bytes)
{
    if (bytes <= *cap) return MANTIS_OK;
    if (*p) { MCHECK(c, cudaFreeHost(*p)); *p = nullptr; *cap = 0; }
    MCHECK(c, cudaHostAlloc(p, bytes, cudaHostAllocDefault));
    *cap = bytes;
    return MANTIS_OK;
}

// ---- stage launchers (defined in the .cu files) ----
// Candidate kernel: BGR -> gray -> blur -> Sobel -> NMS -> weak / strong bit planes.
int launch_edge_candidates(mantis_ctx* c, const uint8_t* d_bgr, int n, int w, int h, int stride,
                           size_t frame_pitch, int low, int high, uint32_t* d_weak, uint32_t* d_strong);
// The same stage in marching form (edge_march.cu) for 16-byte aligned rows and widths that are multiples of 16.
bool edge_march_supported(const uint8_t* d_bgr, int w, int h, int stride, size_t frame_pitch);
int launch_edge_march(mantis_ctx* c, const uint8_t* d_bgr, int n, int w, int h, int stride, size_t frame_pitch, int low,
                      int high, uint32_t* d_weak, uint32_t* d_strong);
// Per-frame on-chip kernel: hysteresis closure, dilate x2 + erode x1, edge mask; every output optional.
struct EdgeOutputs {
    uint32_t* canny_bits = nullptr;
    uint32_t* morph_bits = nullptr;
    uint32_t* mask_bits = nullptr;
    uint8_t* canny_u8 = nullptr;
    uint8_t* morph_u8 = nullptr;
    uint8_t* mask_u8 = nullptr;
};
int launch_edge_link(mantis_ctx* c, const uint32_t* d_weak, const uint32_t* d_strong, int n, int w, int h,
                     const EdgeOutputs& out);

// Contours -> approxPolyDP -> 4-gons -> de-duplicated quads, per frame, in OpenCV's order.
int launch_quads(mantis_ctx* c, const uint32_t* d_morph_bits, int n, int w, int h, mantis_frame_quads* d_out);

namespace pm { struct Camera; }
// cams_per_unit > 1 (or d_b2c_inv != null): the frames are rigs of cams_per_unit cameras whose hypotheses are pooled
// in the base frame (per-frame d_b2c_inv); n must be a multiple of cams_per_unit.
int launch_pose(mantis_ctx* c, const FrameQuads* d_quads, const pm::Camera* d_cams, int n, Hyp* d_hyps, short* d_sel,
                FramePose* d_info, int cams_per_unit = 1, const Tf* d_b2c_inv = nullptr);
int launch_search(mantis_ctx* c, const uint8_t* d_bgr, size_t pitch, int stride, const uint32_t* d_mask_bits, int n,
                  int w, int h, const pm::Camera* d_cams, const FrameQuads* d_quads, const Hyp* d_hyps,
                  const short* d_sel, const FramePose* d_info, mantis_result* d_out, mantis_frame_trace* d_trace);
int launch_rpp_planar4(mantis_ctx* c, const double* d_model_xyz, int model_shared, const double* d_image_xy, int n,
                       mantis_rpp_result* d_out);
int launch_rpp_generic(mantis_ctx* c, const double* d_model_xyz, int model_shared, const double* d_image_xy, int n, int m,
                       mantis_rpp_result* d_out);
int launch_search_rig(mantis_ctx* c, const uint8_t* d_bgr, size_t pitch, int stride, const uint32_t* d_mask_bits, int n_rigs,
                      int rig_cams, int w, int h, const pm::Camera* d_cams, const FrameQuads* d_quads, const Hyp* d_hyps,
                      const short* d_sel, const FramePose* d_info, const Tf* d_b2c, mantis_result* d_out,
                      mantis_frame_trace* d_trace);
int launch_score(mantis_ctx* c, const uint8_t* d_bgr, size_t pitch, int stride, const uint32_t* d_mask_bits, int n, int w,
                 int h, const pm::Camera* d_cams, const double* d_w2c, int n_hyps, double* d_scores);
extern "C" int run_edge_stage(mantis_ctx* c, const uint8_t* d_in, int n, int width, int height, int stride_bytes,
                   size_t frame_pitch_bytes, const EdgeOutputs& out);

// argument validation / host->device staging shared by the stage entry points (api.cu)
int check_frames(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                 size_t frame_pitch_bytes, int mem);
int stage_input(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                size_t frame_pitch_bytes, int mem, const uint8_t** d_in);

static inline int words_per_row(int w) { return (w + 31) >> 5; }
