// api.cu -- C ABI entry points (include/mantis_cuda.h): context life cycle and the
// image-stage batch call.  No CPU fallback: without a CUDA device creation fails.
#include <new>
#include "ctx.h"

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

int mantis_create(const mantis_config* cfg, int device, mantis_ctx** out)
{
    if (!out) return MANTIS_ERR_INVALID;
    *out = nullptr;
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
    *out = c;
    return MANTIS_OK;
}

int mantis_destroy(mantis_ctx* c)
{
    if (!c) return MANTIS_ERR_INVALID;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    DevBuf* bufs[] = {&c->weak, &c->strong, &c->canny_bits, &c->morph_bits, &c->stage_in, &c->stage_out0,
                      &c->stage_out1};
    for (DevBuf* b : bufs)
        if (b->p) cudaFree(b->p);
    if (c->pin_in) cudaFreeHost(c->pin_in);
    if (c->pin_out) cudaFreeHost(c->pin_out);
    cudaStreamDestroy(c->stream);
    delete c;
    return MANTIS_OK;
}

const char* mantis_last_error(const mantis_ctx* c) { return c ? c->err : ""; }
uint64_t mantis_launch_count(const mantis_ctx* c) { return c ? c->launches : 0; }
void* mantis_stream(const mantis_ctx* c) { return c ? (void*)c->stream : nullptr; }

int mantis_edges_batch(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                       size_t frame_pitch_bytes, uint8_t* canny_out, uint8_t* morph_out, int mem)
{
    if (!c) return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    if (!bgr || n < 0 || width <= 0 || height <= 0 || stride_bytes < 3 * width ||
        frame_pitch_bytes < (size_t)stride_bytes * (height - 1) + (size_t)3 * width ||
        (mem != MANTIS_MEM_HOST && mem != MANTIS_MEM_DEVICE))
        return MANTIS_ERR_INVALID;
    MCHECK(c, cudaSetDevice(c->device));

    const int WW = words_per_row(width);
    const size_t plane_bytes = (size_t)n * height * WW * sizeof(uint32_t);
    const size_t map_bytes = (size_t)n * height * width;
    int rc;
    if ((rc = ctx_reserve(c, c->weak, plane_bytes)) || (rc = ctx_reserve(c, c->strong, plane_bytes))) return rc;

    const uint8_t* d_in = bgr;
    uint8_t* d_canny = canny_out;
    uint8_t* d_morph = morph_out;
    if (mem == MANTIS_MEM_HOST) {
        // stage the whole batch: H2D on the context stream, results D2H after the kernels
        size_t in_bytes = frame_pitch_bytes * (size_t)(n - 1) + (size_t)stride_bytes * (height - 1) + 3 * (size_t)width;
        if ((rc = ctx_reserve(c, c->stage_in, in_bytes))) return rc;
        MCHECK(c, cudaMemcpyAsync(c->stage_in.p, bgr, in_bytes, cudaMemcpyHostToDevice, c->stream));
        d_in = (const uint8_t*)c->stage_in.p;
        if (canny_out) { if ((rc = ctx_reserve(c, c->stage_out0, map_bytes))) return rc; d_canny = (uint8_t*)c->stage_out0.p; }
        if (morph_out) { if ((rc = ctx_reserve(c, c->stage_out1, map_bytes))) return rc; d_morph = (uint8_t*)c->stage_out1.p; }
    }

    const int low = c->cfg.canny_hysteresis, high = 3 * c->cfg.canny_hysteresis;
    // gridDim.z is limited to 65535 frames per launch
    for (int f0 = 0; f0 < n; f0 += 32768) {
        int nf = (n - f0 < 32768) ? (n - f0) : 32768;
        size_t poff = (size_t)f0 * height * WW;
        if ((rc = launch_edge_candidates(c, d_in + (size_t)f0 * frame_pitch_bytes, nf, width, height, stride_bytes,
                                         frame_pitch_bytes, low, high, (uint32_t*)c->weak.p + poff,
                                         (uint32_t*)c->strong.p + poff)))
            return rc;
        size_t moff = (size_t)f0 * height * width;
        if ((rc = launch_edge_link(c, (uint32_t*)c->weak.p + poff, (uint32_t*)c->strong.p + poff, nf, width, height,
                                   nullptr, nullptr, d_canny ? d_canny + moff : nullptr,
                                   d_morph ? d_morph + moff : nullptr)))
            return rc;
    }
    if (mem == MANTIS_MEM_HOST) {
        if (canny_out) MCHECK(c, cudaMemcpyAsync(canny_out, d_canny, map_bytes, cudaMemcpyDeviceToHost, c->stream));
        if (morph_out) MCHECK(c, cudaMemcpyAsync(morph_out, d_morph, map_bytes, cudaMemcpyDeviceToHost, c->stream));
    }
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

}  // extern "C"
