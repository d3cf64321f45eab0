// noalloc.cu -- mantis_edges_noalloc(): the image stage (K1 + K2, same kernels as mantis_edges_batch) for
// callers that cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses
// that forbid it.  No counterpart in the reference.
//
// mantis_edges_batch keeps K1's output (the weak / strong candidate bit planes that K2 reads back) in
// context scratch that grows with cudaMalloc.  Here that scratch is a file-scope __device__ array of a
// fixed size, so the call allocates nothing:
//   - every GPU that loads the library holds the array (MANTIS_NOALLOC_SCRATCH_BYTES = 8 MiB) until exit;
//   - a batch whose two bit planes do not fit is rejected with MANTIS_ERR_INVALID and a message in
//     mantis_last_error() -- nothing is truncated;
//   - all calls on a device share the array and it keeps the last call's contents: the region a call uses
//     is cleared (cudaMemsetAsync, on the context's stream) before K1 writes it, and two calls that overlap
//     on the GPU (two contexts / two streams on one device) RACE on it.  Serialise them.
// In host code the array's name is a host-side stand-in, not a device address; the device address is taken
// once per device with cudaGetSymbolAddress (that device current) and only that pointer goes to the kernels.
#include <mutex>
#include "ctx.h"

namespace {

constexpr size_t SCRATCH_BYTES = MANTIS_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned char* g_scratch_addr[MAX_DEVICES] = {};   // device address of g_noalloc_scratch, per device

// Device address of the scratch array on `device` (which must be the current device).
int scratch_address(mantis_ctx* c, int device, unsigned char** out)
{
    if (device < 0 || device >= MAX_DEVICES) {
        snprintf(c->err, sizeof(c->err), "mantis_edges_noalloc: device %d beyond the %d devices supported", device, MAX_DEVICES);
        return MANTIS_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[device]) {
        unsigned char* p = nullptr;
        MCHECK(c, cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) {
            snprintf(c->err, sizeof(c->err), "mantis_edges_noalloc: cudaGetSymbolAddress returned a null address");
            return MANTIS_ERR_CUDA;
        }
        g_scratch_addr[device] = p;
    }
    *out = g_scratch_addr[device];
    return MANTIS_OK;
}

}  // namespace

extern "C" int mantis_edges_noalloc(mantis_ctx* c, const uint8_t* d_bgr, int n, int width, int height, int stride_bytes,
                                    size_t frame_pitch_bytes, uint8_t* d_canny_out, uint8_t* d_morph_out)
{
    int rc = check_frames(c, d_bgr, n, width, height, stride_bytes, frame_pitch_bytes, MANTIS_MEM_DEVICE);
    if (rc) return rc;
    if (!d_canny_out && !d_morph_out) return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    const int WW = words_per_row(width);
    const size_t plane_bytes = (size_t)n * height * WW * sizeof(uint32_t);   // one bit plane of the batch
    const size_t plane_pitch = (plane_bytes + 255) & ~(size_t)255;           // keeps the second plane 256-byte aligned
    if (n > 32768 || plane_pitch > SCRATCH_BYTES / 2) {
        snprintf(c->err, sizeof(c->err),
                 "mantis_edges_noalloc: %d frames of %dx%d need 2 x %zu bytes of bit planes, the fixed scratch holds 2 x %zu "
                 "(MANTIS_NOALLOC_SCRATCH_BYTES = %zu); split the batch or use mantis_edges_batch",
                 n, width, height, plane_bytes, SCRATCH_BYTES / 2, SCRATCH_BYTES);
        return MANTIS_ERR_INVALID;
    }
    MCHECK(c, cudaSetDevice(c->device));
    unsigned char* base = nullptr;
    if ((rc = scratch_address(c, c->device, &base))) return rc;
    uint32_t* d_weak = (uint32_t*)base;
    uint32_t* d_strong = (uint32_t*)(base + plane_pitch);
    // the array keeps the previous call's contents: what this call uses starts at zero
    MCHECK(c, cudaMemsetAsync(base, 0, 2 * plane_pitch, c->stream));
    const int low = c->cfg.canny_hysteresis, high = 3 * c->cfg.canny_hysteresis;
    if (edge_march_supported(d_bgr, width, height, stride_bytes, frame_pitch_bytes))
        rc = launch_edge_march(c, d_bgr, n, width, height, stride_bytes, frame_pitch_bytes, low, high, d_weak, d_strong);
    else
        rc = launch_edge_candidates(c, d_bgr, n, width, height, stride_bytes, frame_pitch_bytes, low, high, d_weak, d_strong);
    if (rc) return rc;
    EdgeOutputs eo;
    eo.canny_u8 = d_canny_out;
    eo.morph_u8 = d_morph_out;
    if ((rc = launch_edge_link(c, d_weak, d_strong, n, width, height, eo))) return rc;
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}
