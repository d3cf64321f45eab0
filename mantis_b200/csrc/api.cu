// api.cu -- C ABI entry points (include/mantis_cuda.h): context life cycle and the
// image-stage batch call.  No CPU fallback: without a CUDA device creation fails.
#include <new>
#include <cmath>
#include <stdlib.h>
#include "ctx.h"

int ctx_init_tables(mantis_ctx* c);

extern "C" {

void mantis_default_config(mantis_config* cfg)
{
    if (!cfg) return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->canny_hysteresis = 50;
    cfg->polygon_epsilon = 10;
    cfg->search_radius_multiplier = 0.1;
    cfg->grid_spacing = 0.32;
    cfg->grid_size = 9;
    cfg->quad_stretch_factor = 1.3;
    cfg->max_quad_error = 0.5;
    cfg->max_angle_difference = 0.2;
    cfg->projection_bias = 1.1;
    cfg->color_projection_bias = 1.1;
    cfg->rot_sigma = 0.03;
    cfg->trans_sigma = 0.01;
    cfg->particles = 50;
    cfg->iterations = 10;
    cfg->rng_seed = 1;
}

const char* mantis_strerror(int status)
{
    switch (status) {
        case MANTIS_OK: return "ok";
        case MANTIS_ERR_INVALID: return "invalid argument";
        case MANTIS_ERR_CUDA: return "CUDA runtime error";
        case MANTIS_ERR_NO_DEVICE: return "no CUDA device (libmantis_cuda has no CPU fallback)";
        case MANTIS_ERR_UNSUPPORTED: return "unsupported frame size";
        case MANTIS_ERR_NOMEM: return "out of memory";
        default: return "unknown status";
    }
}

// Limits the kernels are built for (search.cu: 96 particles / 96 cell shifts per frame), checked once here
// instead of on the first localize call.
static bool config_valid(const mantis_config& g)
{
    auto pos = [](double v) { return std::isfinite(v) && v > 0; };
    if (g.canny_hysteresis < 0 || g.canny_hysteresis > 680 || g.polygon_epsilon < 0) return false;   // 3T must stay below the fp16 magnitude range (2040)
    if (g.particles < 0 || g.particles > 96 || g.iterations < 0 || g.iterations > 4096) return false;
    if ((long long)g.particles * g.iterations > 100000) return false;
    if (g.grid_size < 1 || (long long)g.grid_size * g.grid_size > 96) return false;
    if (!pos(g.grid_spacing) || !pos(g.projection_bias) || !pos(g.color_projection_bias)) return false;
    if (!pos(g.quad_stretch_factor) || !std::isfinite(g.search_radius_multiplier) || g.search_radius_multiplier < 0) return false;
    if (!std::isfinite(g.max_quad_error) || !std::isfinite(g.max_angle_difference) || g.max_angle_difference < 0) return false;
    if (!std::isfinite(g.rot_sigma) || g.rot_sigma < 0 || !std::isfinite(g.trans_sigma) || g.trans_sigma < 0) return false;
    return true;
}

int mantis_create(const mantis_config* cfg, int device, mantis_ctx** out)
{
    if (!out) return MANTIS_ERR_INVALID;
    *out = nullptr;
    if (cfg && !config_valid(*cfg)) return MANTIS_ERR_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MANTIS_ERR_NO_DEVICE;
    if (device < 0 || device >= ndev) return MANTIS_ERR_INVALID;
    mantis_ctx* c = new (std::nothrow) mantis_ctx();
    if (!c) return MANTIS_ERR_NOMEM;
    if (cfg) c->cfg = *cfg; else mantis_default_config(&c->cfg);
    c->device = device;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) {
        delete c;
        return MANTIS_ERR_CUDA;
    }
    int rc = ctx_init_tables(c);
    if (rc != MANTIS_OK) {
        mantis_destroy(c);
        return rc;
    }
    *out = c;
    return MANTIS_OK;
}

int mantis_destroy(mantis_ctx* c)
{
    if (!c) return MANTIS_ERR_INVALID;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    for (int i = 0; i < mantis_ctx::MAX_LANES; ++i)
        if (c->lane[i]) { mantis_destroy(c->lane[i]); c->lane[i] = nullptr; }
    DevBuf* bufs[] = {&c->weak, &c->strong, &c->canny_bits, &c->morph_bits, &c->stage_in, &c->stage_in2, &c->stage_out0,
                      &c->stage_out1, &c->q_pts, &c->q_dst, &c->q_stack, &c->q_owner, &c->q_pad, &c->q_tmp, &c->frame_quads,
                      &c->cams, &c->hyps, &c->sel, &c->pose_info, &c->results, &c->traces, &c->mask_bits,
                      &c->d_map, &c->d_rand, &c->d_shifts, &c->d_shift_pts, &c->d_shift_idx, &c->render_poses, &c->quad_hyp, &c->pose_offs,
                      &c->hough_trig, &c->hough_count, &c->hough_peaks, &c->hough_lines, &c->d_atan, &c->search_ticket, &c->rig_tf, &c->pnp_stage, &c->leg_a, &c->leg_b, &c->leg_c};
    for (DevBuf* b : bufs)
        if (b->p) cudaFree(b->p);
    for (int i = 0; i <= MANTIS_N_STAGES; ++i)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    for (int i = 0; i < 2; ++i) {
        if (c->ev_copied[i]) cudaEventDestroy(c->ev_copied[i]);
        if (c->ev_free[i]) cudaEventDestroy(c->ev_free[i]);
    }
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    cudaStreamDestroy(c->stream);
    delete c;
    return MANTIS_OK;
}

// Tuning aid (not part of the public header): accumulated clock64 deltas of the quad kernel's phases.
int mantis_debug_clocks(mantis_ctx* c, long long* out16)
{
    if (!c) return MANTIS_ERR_INVALID;
    MCHECK(c, cudaSetDevice(c->device));
    if (!c->dbg_clocks) {
        MCHECK(c, cudaMalloc(&c->dbg_clocks, 16 * sizeof(long long)));
        MCHECK(c, cudaMemset(c->dbg_clocks, 0, 16 * sizeof(long long)));
    }
    if (out16) {
        MCHECK(c, cudaMemcpy(out16, c->dbg_clocks, 16 * sizeof(long long), cudaMemcpyDeviceToHost));
        MCHECK(c, cudaMemset(c->dbg_clocks, 0, 16 * sizeof(long long)));
    }
    return MANTIS_OK;
}

int mantis_set_profiling(mantis_ctx* c, int enabled)
{
    if (!c) return MANTIS_ERR_INVALID;
    MCHECK(c, cudaSetDevice(c->device));
    if (enabled && !c->ev[0])
        for (int i = 0; i <= MANTIS_N_STAGES; ++i) MCHECK(c, cudaEventCreate(&c->ev[i]));
    c->profiling = enabled != 0;
    for (int i = 0; i < MANTIS_N_STAGES; ++i) c->stage_ms[i] = 0;
    return MANTIS_OK;
}

int mantis_get_stage_ms(mantis_ctx* c, double* ms)
{
    if (!c || !ms) return MANTIS_ERR_INVALID;
    for (int i = 0; i < MANTIS_N_STAGES; ++i) { ms[i] = c->stage_ms[i]; c->stage_ms[i] = 0; }
    return MANTIS_OK;
}

const char* mantis_last_error(const mantis_ctx* c) { return c ? c->err : ""; }
uint64_t mantis_launch_count(const mantis_ctx* c)
{
    if (!c) return 0;
    uint64_t n = c->launches;
    for (int i = 0; i < mantis_ctx::MAX_LANES; ++i)
        if (c->lane[i]) n += c->lane[i]->launches;
    return n;
}
void* mantis_stream(const mantis_ctx* c) { return c ? (void*)c->stream : nullptr; }

}  // extern "C"

// ---- shared plumbing --------------------------------------------------------------------
int check_frames(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                        size_t frame_pitch_bytes, int mem)
{
    if (!c) return MANTIS_ERR_INVALID;
    if (!bgr || n < 0 || width <= 0 || height <= 0 || width > 32767 || height > 32767 || stride_bytes < 3 * width ||
        frame_pitch_bytes < (size_t)stride_bytes * (height - 1) + (size_t)3 * width ||
        (mem != MANTIS_MEM_HOST && mem != MANTIS_MEM_DEVICE))
        return MANTIS_ERR_INVALID;
    return MANTIS_OK;
}

// Brings the frames onto the device if the caller holds them on the host.
int stage_input(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                       size_t frame_pitch_bytes, int mem, const uint8_t** d_in)
{
    *d_in = bgr;
    if (mem == MANTIS_MEM_HOST) {
        size_t in_bytes = frame_pitch_bytes * (size_t)(n - 1) + (size_t)stride_bytes * (height - 1) + 3 * (size_t)width;
        int rc;
        if ((rc = ctx_reserve(c, c->stage_in, in_bytes))) return rc;
        MCHECK(c, cudaMemcpyAsync(c->stage_in.p, bgr, in_bytes, cudaMemcpyHostToDevice, c->stream));
        *d_in = (const uint8_t*)c->stage_in.p;
    }
    return MANTIS_OK;
}

// K1 + K2 over a batch already on the device.  `out` pointers are per-batch bases.
int run_edge_stage(mantis_ctx* c, const uint8_t* d_in, int n, int width, int height, int stride_bytes,
                   size_t frame_pitch_bytes, const EdgeOutputs& out)
{
    const int WW = words_per_row(width);
    const size_t plane_bytes = (size_t)n * height * WW * sizeof(uint32_t);
    int rc;
    if ((rc = ctx_reserve(c, c->weak, plane_bytes)) || (rc = ctx_reserve(c, c->strong, plane_bytes))) return rc;
    const int low = c->cfg.canny_hysteresis, high = 3 * c->cfg.canny_hysteresis;
    // gridDim.z is limited to 65535 frames per launch
    for (int f0 = 0; f0 < n; f0 += 32768) {
        int nf = (n - f0 < 32768) ? (n - f0) : 32768;
        size_t poff = (size_t)f0 * height * WW, moff = (size_t)f0 * height * width;
        const uint8_t* d_f0 = d_in + (size_t)f0 * frame_pitch_bytes;
        static const bool force_tiled = getenv("MANTIS_K1_TILED") != nullptr;   // tuning aid: the tiled kernel for every shape
        if (!force_tiled && edge_march_supported(d_f0, width, height, stride_bytes, frame_pitch_bytes))
            rc = launch_edge_march(c, d_f0, nf, width, height, stride_bytes, frame_pitch_bytes, low, high,
                                   (uint32_t*)c->weak.p + poff, (uint32_t*)c->strong.p + poff);
        else
            rc = launch_edge_candidates(c, d_f0, nf, width, height, stride_bytes, frame_pitch_bytes, low, high,
                                        (uint32_t*)c->weak.p + poff, (uint32_t*)c->strong.p + poff);
        if (rc) return rc;
        prof_mark(c, 2);
        EdgeOutputs eo;
        eo.canny_bits = out.canny_bits ? out.canny_bits + poff : nullptr;
        eo.morph_bits = out.morph_bits ? out.morph_bits + poff : nullptr;
        eo.mask_bits = out.mask_bits ? out.mask_bits + poff : nullptr;
        eo.canny_u8 = out.canny_u8 ? out.canny_u8 + moff : nullptr;
        eo.morph_u8 = out.morph_u8 ? out.morph_u8 + moff : nullptr;
        eo.mask_u8 = out.mask_u8 ? out.mask_u8 + moff : nullptr;
        if ((rc = launch_edge_link(c, (uint32_t*)c->weak.p + poff, (uint32_t*)c->strong.p + poff, nf, width, height, eo)))
            return rc;
    }
    return MANTIS_OK;
}

extern "C" {

int mantis_edges_batch(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                       size_t frame_pitch_bytes, uint8_t* canny_out, uint8_t* morph_out, int mem)
{
    int rc = check_frames(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem);
    if (rc) return rc;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    const size_t map_bytes = (size_t)n * height * width;
    const uint8_t* d_in;
    if ((rc = stage_input(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem, &d_in))) return rc;
    EdgeOutputs eo;
    eo.canny_u8 = canny_out;
    eo.morph_u8 = morph_out;
    if (mem == MANTIS_MEM_HOST) {
        if (canny_out) { if ((rc = ctx_reserve(c, c->stage_out0, map_bytes))) return rc; eo.canny_u8 = (uint8_t*)c->stage_out0.p; }
        if (morph_out) { if ((rc = ctx_reserve(c, c->stage_out1, map_bytes))) return rc; eo.morph_u8 = (uint8_t*)c->stage_out1.p; }
    }
    if ((rc = run_edge_stage(c, d_in, n, width, height, stride_bytes, frame_pitch_bytes, eo))) return rc;
    if (mem == MANTIS_MEM_HOST) {
        if (canny_out) MCHECK(c, cudaMemcpyAsync(canny_out, eo.canny_u8, map_bytes, cudaMemcpyDeviceToHost, c->stream));
        if (morph_out) MCHECK(c, cudaMemcpyAsync(morph_out, eo.morph_u8, map_bytes, cudaMemcpyDeviceToHost, c->stream));
    }
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

int mantis_mask_batch(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                      size_t frame_pitch_bytes, uint8_t* mask_out, int mem)
{
    int rc = check_frames(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem);
    if (rc) return rc;
    if (!mask_out) return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    const uint8_t* d_in;
    if ((rc = stage_input(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem, &d_in))) return rc;
    const size_t plane_bytes = (size_t)n * height * words_per_row(width) * sizeof(uint32_t);
    const size_t map_bytes = (size_t)n * height * width;
    if ((rc = ctx_reserve(c, c->canny_bits, plane_bytes))) return rc;
    EdgeOutputs eo;
    eo.canny_bits = (uint32_t*)c->canny_bits.p;
    eo.mask_u8 = mask_out;
    if (mem == MANTIS_MEM_HOST) {
        if ((rc = ctx_reserve(c, c->stage_out0, map_bytes))) return rc;
        eo.mask_u8 = (uint8_t*)c->stage_out0.p;
    }
    if ((rc = run_edge_stage(c, d_in, n, width, height, stride_bytes, frame_pitch_bytes, eo))) return rc;
    if (mem == MANTIS_MEM_HOST)
        MCHECK(c, cudaMemcpyAsync(mask_out, eo.mask_u8, map_bytes, cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

int mantis_quads_batch(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                       size_t frame_pitch_bytes, mantis_frame_quads* out, int mem)
{
    int rc = check_frames(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem);
    if (rc) return rc;
    if (!out) return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    const uint8_t* d_in;
    if ((rc = stage_input(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem, &d_in))) return rc;
    const size_t plane_bytes = (size_t)n * height * words_per_row(width) * sizeof(uint32_t);
    if ((rc = ctx_reserve(c, c->morph_bits, plane_bytes))) return rc;
    EdgeOutputs eo;
    eo.morph_bits = (uint32_t*)c->morph_bits.p;
    if ((rc = run_edge_stage(c, d_in, n, width, height, stride_bytes, frame_pitch_bytes, eo))) return rc;
    mantis_frame_quads* d_out = out;
    if (mem == MANTIS_MEM_HOST) {
        if ((rc = ctx_reserve(c, c->frame_quads, (size_t)n * sizeof(mantis_frame_quads)))) return rc;
        d_out = (mantis_frame_quads*)c->frame_quads.p;
    }
    if ((rc = launch_quads(c, eo.morph_bits, n, width, height, d_out))) return rc;
    if (mem == MANTIS_MEM_HOST)
        MCHECK(c, cudaMemcpyAsync(out, d_out, (size_t)n * sizeof(mantis_frame_quads), cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

}  // extern "C"
