// render.cu -- device-side synthetic frame source: the IARC floor grid of params/map.yaml seen by
// a fisheye (Kannala-Brandt) camera.  Stands in for the reference's OpenGL GridRenderer
// (include/mantis3/GridRenderer.cpp:211-353; test/unit_test.cpp:224-233), whose GL path cannot run
// here: one ray per pixel centre, cast onto the plane z = 0.  Unlike the reference's renderer
// (identity projection, GridRenderer.cpp:306-317; pitch 0.3, :232) it uses the camera intrinsics
// and the map's pitch.  Optional additive noise (counter-based hash, N(0, sigma) by Box-Muller)
// reproduces BASELINE config 4's "noise" frames.
#include "ctx.h"
#include "posemath.cuh"

namespace {

struct RenderArgs {
    const double* poses;  // per frame: R_w2c (9, row-major) + t_w2c (3)
    pm::Camera cam;
    int W, H, n;
    double line_width, half, pitch;
    int n_lines, floor_level;
    float noise_sigma;
    uint32_t seed;
    uint8_t* out;
    size_t frame_pitch;
    int stride;
};

__device__ __forceinline__ uint32_t hash32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

__global__ void __launch_bounds__(256) k_render(RenderArgs a)
{
    const int u = blockIdx.x * 64 + (threadIdx.x & 63), v = blockIdx.y * 4 + (threadIdx.x >> 6), f = blockIdx.z;
    if (u >= a.W || v >= a.H) return;
    const double* P = a.poses + 12 * (size_t)f;
    int b = a.floor_level, g = a.floor_level, r = a.floor_level;
    double xn, yn;
    if (pm::fisheye_undistort(a.cam, (double)u, (double)v, xn, yn)) {
        double dw[3];
        for (int i = 0; i < 3; ++i) dw[i] = P[i * 3] * xn + P[i * 3 + 1] * yn + P[i * 3 + 2];
        const double tz = P[11];
        if (dw[2] < 0 && tz > 0) {
            double s = -tz / dw[2];
            double X = P[9] + s * dw[0], Y = P[10] + s * dw[1];
            const double hw = a.line_width / 2;
            if (fabs(X) <= a.half + hw && fabs(Y) <= a.half + hw) {
                double kx = floor((X + a.half) / a.pitch + 0.5), ky = floor((Y + a.half) / a.pitch + 0.5);
                double lx = -a.half + kx * a.pitch, ly = -a.half + ky * a.pitch;
                bool on_x = fabs(X - lx) <= hw, on_y = fabs(Y - ly) <= hw;
                if (on_x || on_y) {
                    if (on_y && ky == (double)(a.n_lines - 1)) { b = 50; g = 255; r = 85; }   // green edge
                    else if (on_y && ky == 0.0) { b = 50; g = 85; r = 255; }                  // red edge
                    else { b = g = r = 255; }
                }
            }
        }
    }
    if (a.noise_sigma > 0.f) {
        uint32_t idx = ((uint32_t)f * (uint32_t)a.H + (uint32_t)v) * (uint32_t)a.W + (uint32_t)u;
        int ch[3] = {b, g, r};
        for (int k = 0; k < 3; ++k) {
            uint32_t h1 = hash32(idx * 3u + k + a.seed), h2 = hash32(h1 ^ 0x9e3779b9u);
            float u1 = ((h1 >> 8) + 1) * (1.0f / 16777216.0f), u2 = (h2 >> 8) * (1.0f / 16777216.0f);
            float nrm = sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
            ch[k] = min(255, max(0, __float2int_rn((float)ch[k] + a.noise_sigma * nrm)));
        }
        b = ch[0]; g = ch[1]; r = ch[2];
    }
    uint8_t* o = a.out + (size_t)f * a.frame_pitch + (size_t)v * a.stride + 3 * u;
    o[0] = (uint8_t)b; o[1] = (uint8_t)g; o[2] = (uint8_t)r;
}

// BASELINE configs[3]'s degradation: Gaussian blur (sigma px, 7x7 taps, BORDER_REFLECT_101) then additive
// N(0, noise_sigma) per channel from the same counter-based hash, rounded and clipped to u8.
struct DegradeArgs {
    const uint8_t* in;
    uint8_t* out;
    int W, H, stride_in, stride_out;
    size_t pitch_in, pitch_out;
    float taps[7];
    float noise_sigma;
    uint32_t seed;
};

__global__ void __launch_bounds__(256) k_degrade(DegradeArgs a)
{
    const int u = blockIdx.x * 64 + (threadIdx.x & 63), v = blockIdx.y * 4 + (threadIdx.x >> 6), f = blockIdx.z;
    if (u >= a.W || v >= a.H) return;
    const uint8_t* src = a.in + (size_t)f * a.pitch_in;
    float acc[3] = {0.f, 0.f, 0.f};
    for (int dy = -3; dy <= 3; ++dy) {
        int y = v + dy;
        y = y < 0 ? -y : (y >= a.H ? 2 * a.H - 2 - y : y);
        y = min(max(y, 0), a.H - 1);
        float row[3] = {0.f, 0.f, 0.f};
        for (int dx = -3; dx <= 3; ++dx) {
            int x = u + dx;
            x = x < 0 ? -x : (x >= a.W ? 2 * a.W - 2 - x : x);
            x = min(max(x, 0), a.W - 1);
            const uint8_t* p = src + (size_t)y * a.stride_in + 3 * x;
            const float t = a.taps[dx + 3];
            row[0] += t * (float)p[0]; row[1] += t * (float)p[1]; row[2] += t * (float)p[2];
        }
        const float t = a.taps[dy + 3];
        acc[0] += t * row[0]; acc[1] += t * row[1]; acc[2] += t * row[2];
    }
    const uint32_t idx = ((uint32_t)f * (uint32_t)a.H + (uint32_t)v) * (uint32_t)a.W + (uint32_t)u;
    uint8_t* o = a.out + (size_t)f * a.pitch_out + (size_t)v * a.stride_out + 3 * u;
    for (int k = 0; k < 3; ++k) {
        float val = acc[k];
        if (a.noise_sigma > 0.f) {
            uint32_t h1 = hash32(idx * 3u + k + a.seed), h2 = hash32(h1 ^ 0x9e3779b9u);
            float u1 = ((h1 >> 8) + 1) * (1.0f / 16777216.0f), u2 = (h2 >> 8) * (1.0f / 16777216.0f);
            val += a.noise_sigma * sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
        }
        o[k] = (uint8_t)min(255, max(0, __float2int_rn(val)));
    }
}

}  // namespace

extern "C" int mantis_degrade_batch(mantis_ctx* c, const uint8_t* d_bgr_in, uint8_t* d_bgr_out, int n, int width, int height,
                                    int stride_in, size_t pitch_in, int stride_out, size_t pitch_out, float blur_sigma,
                                    float noise_sigma, uint32_t seed)
{
    if (!c || !d_bgr_in || !d_bgr_out || d_bgr_in == d_bgr_out || n < 0 || width <= 0 || height <= 0 ||
        stride_in < 3 * width || stride_out < 3 * width || !(blur_sigma >= 0.f) || !(noise_sigma >= 0.f))
        return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    DegradeArgs a;
    a.in = d_bgr_in; a.out = d_bgr_out; a.W = width; a.H = height;
    a.stride_in = stride_in; a.stride_out = stride_out; a.pitch_in = pitch_in; a.pitch_out = pitch_out;
    double t[7], sum = 0;
    for (int k = 0; k < 7; ++k) { t[k] = blur_sigma > 0.f ? exp(-(k - 3) * (k - 3) / (2.0 * blur_sigma * blur_sigma)) : (k == 3); sum += t[k]; }
    for (int k = 0; k < 7; ++k) a.taps[k] = (float)(t[k] / sum);
    a.noise_sigma = noise_sigma; a.seed = seed;
    for (int f0 = 0; f0 < n; f0 += 32768) {
        const int nf = n - f0 < 32768 ? n - f0 : 32768;
        DegradeArgs b = a;
        b.in = a.in + (size_t)f0 * pitch_in;
        b.out = a.out + (size_t)f0 * pitch_out;
        dim3 grid((width + 63) / 64, (height + 3) / 4, nf);
        k_degrade<<<grid, 256, 0, c->stream>>>(b);
        c->launches++;
        MCHECK(c, cudaGetLastError());
    }
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

extern "C" int mantis_render_grid_batch(mantis_ctx* c, const double* poses_w2c, int n, int width, int height,
                                        const float* K, const double* D, double line_width, int floor_level,
                                        float noise_sigma, uint32_t seed, uint8_t* d_bgr_out, int stride_bytes,
                                        size_t frame_pitch_bytes)
{
    if (!c || !poses_w2c || !K || !D || !d_bgr_out || n < 0 || width <= 0 || height <= 0 || stride_bytes < 3 * width)
        return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    int rc = ctx_reserve(c, c->render_poses, sizeof(double) * 12 * (size_t)n);
    if (rc) return rc;
    MCHECK(c, cudaMemcpyAsync(c->render_poses.p, poses_w2c, sizeof(double) * 12 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    RenderArgs a;
    a.poses = (const double*)c->render_poses.p;
    a.cam.fx = K[0]; a.cam.fy = K[4]; a.cam.cx = K[2]; a.cam.cy = K[5];
    for (int k = 0; k < 4; ++k) a.cam.D[k] = D[k];
    a.W = width; a.H = height; a.n = n;
    a.line_width = line_width;
    a.pitch = c->cfg.grid_spacing;
    a.n_lines = c->cfg.grid_size + 1;
    a.half = c->cfg.grid_spacing * c->cfg.grid_size / 2.0;
    a.floor_level = floor_level;
    a.noise_sigma = noise_sigma; a.seed = seed;
    a.out = d_bgr_out; a.frame_pitch = frame_pitch_bytes; a.stride = stride_bytes;
    for (int f0 = 0; f0 < n; f0 += 32768) {
        int nf = n - f0 < 32768 ? n - f0 : 32768;
        RenderArgs b = a;
        b.poses = a.poses + 12 * (size_t)f0;
        b.out = a.out + (size_t)f0 * frame_pitch_bytes;
        dim3 grid((width + 63) / 64, (height + 3) / 4, nf);
        k_render<<<grid, 256, 0, c->stream>>>(b);
        c->launches++;
        MCHECK(c, cudaGetLastError());
    }
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}
