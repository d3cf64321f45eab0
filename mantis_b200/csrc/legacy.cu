// legacy.cu -- SURVEY 8(f) rows 3 and 4, batched:
//
//  k_particle_weights   MonteCarlo::project2d / computeCameraError   include/legacy/mantis/MonteCarlo.cpp:169-226
//                       (pinhole projection of the white / red / green map on an UNDISTORTED image, colour
//                       SSE against WHITE / RED / GREEN of MonteCarlo.h:53-55, >= 10 projected points else 1e17)
//                       one CTA per (particle, frame), lanes over map points, exact integer sums.
//  k_undistort_map + k_remap_bgr
//                       cv::fisheye::undistortImage(img, K, D, K)      MonteCarlo.cpp:284
//                       = initUndistortRectifyMap (CV_16SC2 maps; the row accumulators _x += iR(0,0) are
//                       sequential in OpenCV, so one thread walks one row) + cv::remap(INTER_LINEAR,
//                       BORDER_CONSTANT) with 5-bit fractions and 15-bit fixed-point weights.
//  k_markov             MarkovModel: updateWeights / convolve / senseFusion / constructor
//                       include/mantis3/Markov.cpp:14-26,76-124,164-205,227-258 -- one CTA per camera stream,
//                       thread i owns bin i and sums its 360 Gaussian terms in the reference's order.
// Compiled with -fmad=false: the float/double mix of project2d decides which pixel is read.
#include <float.h>
#include <math.h>
#include "ctx.h"

namespace {

constexpr int DEG = 360;
constexpr double LPI = 3.14159265358979323846;

// ---------------------------------------------------------------------------------------------------------
struct WeightArgs {
    const uint8_t* bgr; size_t pitch; int stride, W, H;
    const float* K;        // 9 per frame
    const double* w2c;     // [frame][particle][12]  R row-major + t  (world -> camera, as w2c1 of computeWeight)
    const double* map;     // xyz, white | red | green
    int n_white, n_red, n_green, n_particles;
    double* out;           // [frame][particle]
};

__global__ void __launch_bounds__(128) k_particle_weights(WeightArgs a)
{
    __shared__ unsigned long long s_err[4];
    __shared__ int s_cnt[4];
    const int part = blockIdx.x, frame = blockIdx.y, tid = threadIdx.x;
    const uint8_t* img = a.bgr + (size_t)frame * a.pitch;
    const float* K = a.K + 9 * frame;
    const double* T = a.w2c + ((size_t)frame * a.n_particles + part) * 12;
    const int n = a.n_white + a.n_red + a.n_green;
    unsigned long long err = 0;
    int cnt = 0;
    for (int i = tid; i < n; i += 128) {
        const int set = i < a.n_white ? 0 : (i < a.n_white + a.n_red ? 1 : 2);
        const int c0 = set == 0 ? 255 : (set == 1 ? 113 : 50), c1 = set == 0 ? 255 : (set == 1 ? 48 : 100),
                  c2 = set == 0 ? 255 : (set == 1 ? 32 : 37);
        const double* p = a.map + 3 * i;
        double c[3];
        for (int k = 0; k < 3; ++k) c[k] = T[k * 3] * p[0] + T[k * 3 + 1] * p[1] + T[k * 3 + 2] * p[2] + T[9 + k];
        const float x = (float)(((double)K[0] * c[0]) / c[2] + (double)K[2]);
        const float y = (float)(((double)K[4] * c[1]) / c[2] + (double)K[5]);
        if (x > 0 && x < (float)a.W && y > 0 && y < (float)a.H) {
            const long long pix = (long long)__float2int_rn(y) * a.W + __float2int_rn(x);
            int b = 0, g = 0, r = 0;
            if (pix >= 0 && pix < (long long)a.W * a.H) {
                const uint8_t* q = img + (size_t)(pix / a.W) * a.stride + 3 * (pix % a.W);
                b = q[0]; g = q[1]; r = q[2];
            }
            err += (unsigned)((b - c0) * (b - c0) + (g - c1) * (g - c1) + (r - c2) * (r - c2));
            ++cnt;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        err += __shfl_xor_sync(0xffffffffu, err, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    if ((tid & 31) == 0) { s_err[tid >> 5] = err; s_cnt[tid >> 5] = cnt; }
    __syncthreads();
    if (tid == 0) {
        const unsigned long long e = s_err[0] + s_err[1] + s_err[2] + s_err[3];
        const int m = s_cnt[0] + s_cnt[1] + s_cnt[2] + s_cnt[3];
        double error = (double)e;
        if (m < 10) error = 1e17;
        a.out[(size_t)frame * a.n_particles + part] = error / (double)m;
    }
}

// ---------------------------------------------------------------------------------------------------------
struct MapArgs {
    double iR[9], f0, f1, c0, c1, D[4];
    int W, H;
    short2* map1;
    unsigned short* map2;
};

__device__ __forceinline__ int sat_round(double v)
{
    return v >= 2147483647.0 ? 2147483647 : (v <= -2147483648.0 ? (-2147483647 - 1) : __double2int_rn(v));
}

__global__ void k_undistort_map(MapArgs a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.H) return;
    double _x = i * a.iR[1] + a.iR[2], _y = i * a.iR[4] + a.iR[5], _w = i * a.iR[7] + a.iR[8];
    for (int j = 0; j < a.W; ++j) {
        double u, v;
        if (_w <= 0) {
            u = (_x > 0) ? -INFINITY : INFINITY;
            v = (_y > 0) ? -INFINITY : INFINITY;
        } else {
            const double x = _x / _w, y = _y / _w;
            const double r = sqrt(x * x + y * y);
            const double theta = atan(r);
            const double theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
            const double theta_d = theta * (1 + a.D[0] * theta2 + a.D[1] * theta4 + a.D[2] * theta6 + a.D[3] * theta8);
            const double scale = (r == 0) ? 1.0 : theta_d / r;
            u = a.f0 * x * scale + a.c0;
            v = a.f1 * y * scale + a.c1;
        }
        const int iu = sat_round(u * 32), iv = sat_round(v * 32);
        a.map1[(size_t)i * a.W + j] = make_short2((short)(iu >> 5), (short)(iv >> 5));
        a.map2[(size_t)i * a.W + j] = (unsigned short)((iv & 31) * 32 + (iu & 31));
        _x += a.iR[0]; _y += a.iR[3]; _w += a.iR[6];
    }
}

__global__ void __launch_bounds__(256)
k_remap_bgr(const uint8_t* __restrict__ src, size_t src_pitch, int stride, int W, int H, const short2* __restrict__ map1,
            const unsigned short* __restrict__ map2, uint8_t* __restrict__ dst, size_t dst_pitch, int dst_stride)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y, frame = blockIdx.z;
    if (j >= W) return;
    const uint8_t* s = src + (size_t)frame * src_pitch;
    uint8_t* d = dst + (size_t)frame * dst_pitch + (size_t)i * dst_stride + 3 * j;
    const short2 m = map1[(size_t)i * W + j];
    const int fr = map2[(size_t)i * W + j] & 1023, fx = fr & 31, fy = fr >> 5;
    const int w00 = (32 - fx) * (32 - fy) * 32, w01 = fx * (32 - fy) * 32, w10 = (32 - fx) * fy * 32, w11 = fx * fy * 32;
    const int sx = m.x, sy = m.y;
    int acc[3] = {1 << 14, 1 << 14, 1 << 14};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int xx = sx + (t & 1), yy = sy + (t >> 1);
        const int wt = t == 0 ? w00 : (t == 1 ? w01 : (t == 2 ? w10 : w11));
        if (xx >= 0 && xx < W && yy >= 0 && yy < H) {
            const uint8_t* q = s + (size_t)yy * stride + 3 * xx;
            acc[0] += q[0] * wt; acc[1] += q[1] * wt; acc[2] += q[2] * wt;
        }
    }
    d[0] = (uint8_t)(acc[0] >> 15); d[1] = (uint8_t)(acc[1] >> 15); d[2] = (uint8_t)(acc[2] >> 15);
}

// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double markov_weight(double x, double mu, double stddev, double y)
{
    return (y / (stddev * sqrt(2 * LPI))) * exp(-((x - mu) * (x - mu) / (2 * stddev * stddev)));
}

// yaw <- normalize(gaussian smoothing of yaw), Markov.cpp:76-124; tmp is a second 360-entry buffer
__device__ void markov_update_weights(double* yaw, double* tmp, double stddev, double* s_sum)
{
    const int i = threadIdx.x;
    if (i < DEG) {
        double mx = 0.0;
        const int diff = (DEG / 2) + i;
        if (i <= DEG / 2) {
            for (int j = 0; j < diff; ++j) mx += markov_weight((double)i, (double)j, stddev, yaw[j]);
            for (int j = diff; j < DEG; ++j) mx += markov_weight((double)i, (double)(-(DEG - j)), stddev, yaw[j]);
        } else {
            for (int j = i; j < diff; ++j) mx += markov_weight((double)i, (double)j, stddev, yaw[j % DEG]);
            for (int j = (diff % DEG); j < i; ++j) mx += markov_weight((double)i, (double)j, stddev, yaw[j]);
        }
        tmp[i] = mx;
    }
    __syncthreads();
    if (i == 0) {
        double sum = 0.0;
        for (int k = 0; k < DEG; ++k) sum += tmp[k];
        *s_sum = sum;
    }
    __syncthreads();
    if (i < DEG) yaw[i] = tmp[i] / *s_sum;
    __syncthreads();
}

__device__ __forceinline__ int yaw_bin(double yaw_rad)
{
    double y = yaw_rad * 180.0 / LPI;
    if (y < 0) y += DEG;
    const int b = (int)y;
    return b < 0 ? 0 : (b >= DEG ? DEG - 1 : b);
}

__global__ void __launch_bounds__(384) k_markov(double* planes, int op, const double* arg_a, const double* arg_b)
{
    __shared__ double p[DEG], q[DEG], t[DEG];
    __shared__ double s_sum;
    const int s = blockIdx.x, i = threadIdx.x;
    double* plane = planes + (size_t)s * DEG;
    if (i < DEG) p[i] = plane[i];
    __syncthreads();
    if (op == MANTIS_MARKOV_INIT) {
        if (i < DEG) p[i] = (i == yaw_bin(arg_a[s])) ? 1.0 : 0.0;
        __syncthreads();
        markov_update_weights(p, t, 3, &s_sum);
    } else if (op == MANTIS_MARKOV_CONVOLVE) {
        const double dtheta = arg_a[s], dt = arg_b[s];
        int disp = (int)((int)dtheta * 180 / LPI);
        if (i < DEG) {
            if (dtheta > 0) q[(i + disp) % DEG] = p[i];  // aux[i % 360] = p[i - disp] for i in [disp, 360 + disp)
            else q[i] = p[((-disp) + i) % DEG];
        }
        __syncthreads();
        if (i < DEG) p[i] = q[i];
        __syncthreads();
        markov_update_weights(p, t, (1.0 / 3.0) * dt * 11.5 / 30.0, &s_sum);
    } else {  // MANTIS_MARKOV_SENSE
        if (i < DEG) q[i] = (i == yaw_bin(arg_a[s])) ? 1.0 : 0.0;
        __syncthreads();
        markov_update_weights(q, t, 3.5, &s_sum);
        const double big = sqrt(DBL_MAX);
        if (i < DEG) t[i] = (p[i] * big) * (q[i] * big);
        __syncthreads();
        if (i == 0) {
            double sum = 0.0;
            for (int k = 0; k < DEG; ++k) sum += t[k];
            s_sum = sum;
        }
        __syncthreads();
        if (i < DEG) p[i] = t[i] / s_sum;
        __syncthreads();
    }
    if (i < DEG) plane[i] = p[i];
}

__global__ void k_markov_errors(const double* planes, const int* stream_of, const double* yaw, double* err, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    err[i] = err[i] * 1 / (planes[(size_t)stream_of[i] * DEG + yaw_bin(yaw[i])]);
}

// small helper: stage a host array on the device (or pass a device pointer through)
template <class T>
int to_device(mantis_ctx* c, DevBuf& buf, const T* src, size_t count, int mem, const T** out)
{
    *out = src;
    if (mem == MANTIS_MEM_HOST && src) {
        int rc = ctx_reserve(c, buf, count * sizeof(T));
        if (rc) return rc;
        MCHECK(c, cudaMemcpyAsync(buf.p, src, count * sizeof(T), cudaMemcpyHostToDevice, c->stream));
        *out = (const T*)buf.p;
    }
    return MANTIS_OK;
}

}  // namespace

extern "C" {

int mantis_particle_weights_batch(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                                  size_t frame_pitch_bytes, const float* K, const double* w2c, int n_particles,
                                  double* weights, int mem)
{
    int rc = check_frames(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem);
    if (rc) return rc;
    if (!K || !w2c || !weights || n_particles <= 0 || c->n_map <= 0) return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    const uint8_t* d_in;
    if ((rc = stage_input(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem, &d_in))) return rc;
    WeightArgs a;
    a.bgr = d_in; a.pitch = frame_pitch_bytes; a.stride = stride_bytes; a.W = width; a.H = height;
    if ((rc = to_device(c, c->leg_a, K, (size_t)9 * n, mem, &a.K))) return rc;
    if ((rc = to_device(c, c->leg_b, w2c, (size_t)12 * n * n_particles, mem, &a.w2c))) return rc;
    a.map = (const double*)c->d_map.p;
    a.n_green = c->n_green; a.n_red = c->n_red; a.n_white = c->n_map - c->n_green - c->n_red;
    a.n_particles = n_particles;
    a.out = weights;
    if (mem == MANTIS_MEM_HOST) {
        if ((rc = ctx_reserve(c, c->leg_c, sizeof(double) * (size_t)n * n_particles))) return rc;
        a.out = (double*)c->leg_c.p;
    }
    for (int f0 = 0; f0 < n; f0 += 32768) {
        const int nf = n - f0 < 32768 ? n - f0 : 32768;
        WeightArgs b = a;
        b.bgr += (size_t)f0 * a.pitch; b.K += (size_t)9 * f0; b.w2c += (size_t)12 * f0 * n_particles; b.out += (size_t)f0 * n_particles;
        k_particle_weights<<<dim3(n_particles, nf), 128, 0, c->stream>>>(b);
        c->launches++;
        MCHECK(c, cudaGetLastError());
    }
    if (mem == MANTIS_MEM_HOST)
        MCHECK(c, cudaMemcpyAsync(weights, a.out, sizeof(double) * (size_t)n * n_particles, cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

int mantis_undistort_batch(mantis_ctx* c, const uint8_t* bgr, int n, int width, int height, int stride_bytes,
                           size_t frame_pitch_bytes, const float* K, const double* D, uint8_t* out, int mem)
{
    int rc = check_frames(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem);
    if (rc) return rc;
    if (!K || !D || !out || K[0] == 0 || K[4] == 0 || K[3] != 0 || K[6] != 0 || K[7] != 0 || K[8] != 1) return MANTIS_ERR_INVALID;
    if (n == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    const uint8_t* d_in;
    if ((rc = stage_input(c, bgr, n, width, height, stride_bytes, frame_pitch_bytes, mem, &d_in))) return rc;
    const size_t px = (size_t)width * height, tight = 3 * px;
    if ((rc = ctx_reserve(c, c->leg_a, px * sizeof(short2))) || (rc = ctx_reserve(c, c->leg_b, px * sizeof(unsigned short)))) return rc;
    MapArgs m;
    {   // iR = (Knew * I)^-1 with Knew = K (upper triangular), in double from the float camera matrix
        const double fx = K[0], s = K[1], cx = K[2], fy = K[4], cy = K[5];
        const double iR[9] = {1.0 / fx, -s / (fx * fy), (s * cy - cx * fy) / (fx * fy), 0, 1.0 / fy, -cy / fy, 0, 0, 1};
        for (int k = 0; k < 9; ++k) m.iR[k] = iR[k];
        m.f0 = fx; m.f1 = fy; m.c0 = cx; m.c1 = cy;
        for (int k = 0; k < 4; ++k) m.D[k] = D[k];
    }
    m.W = width; m.H = height;
    m.map1 = (short2*)c->leg_a.p; m.map2 = (unsigned short*)c->leg_b.p;
    k_undistort_map<<<(height + 63) / 64, 64, 0, c->stream>>>(m);
    c->launches++;
    MCHECK(c, cudaGetLastError());
    uint8_t* d_out = out;
    if (mem == MANTIS_MEM_HOST) {
        if ((rc = ctx_reserve(c, c->stage_out0, tight * n))) return rc;
        d_out = (uint8_t*)c->stage_out0.p;
    }
    for (int f0 = 0; f0 < n; f0 += 32768) {
        const int nf = n - f0 < 32768 ? n - f0 : 32768;
        k_remap_bgr<<<dim3((width + 255) / 256, height, nf), 256, 0, c->stream>>>(
            d_in + (size_t)f0 * frame_pitch_bytes, frame_pitch_bytes, stride_bytes, width, height, m.map1, m.map2,
            d_out + (size_t)f0 * tight, tight, 3 * width);
        c->launches++;
        MCHECK(c, cudaGetLastError());
    }
    if (mem == MANTIS_MEM_HOST) MCHECK(c, cudaMemcpyAsync(out, d_out, tight * n, cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

int mantis_markov_batch(mantis_ctx* c, double* planes, int n_streams, int op, const double* a, const double* b, int mem)
{
    if (!c || !planes || n_streams < 0 || !a || (op == MANTIS_MARKOV_CONVOLVE && !b) ||
        (op != MANTIS_MARKOV_INIT && op != MANTIS_MARKOV_CONVOLVE && op != MANTIS_MARKOV_SENSE) ||
        (mem != MANTIS_MEM_HOST && mem != MANTIS_MEM_DEVICE))
        return MANTIS_ERR_INVALID;
    if (n_streams == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    double* d_planes = planes;
    const double *d_a, *d_b = b;
    int rc;
    if (mem == MANTIS_MEM_HOST) {
        if ((rc = ctx_reserve(c, c->leg_c, sizeof(double) * DEG * (size_t)n_streams))) return rc;
        d_planes = (double*)c->leg_c.p;
        MCHECK(c, cudaMemcpyAsync(d_planes, planes, sizeof(double) * DEG * (size_t)n_streams, cudaMemcpyHostToDevice, c->stream));
    }
    if ((rc = to_device(c, c->leg_a, a, (size_t)n_streams, mem, &d_a))) return rc;
    if (b && (rc = to_device(c, c->leg_b, b, (size_t)n_streams, mem, &d_b))) return rc;
    k_markov<<<n_streams, 384, 0, c->stream>>>(d_planes, op, d_a, d_b);
    c->launches++;
    MCHECK(c, cudaGetLastError());
    if (mem == MANTIS_MEM_HOST)
        MCHECK(c, cudaMemcpyAsync(planes, d_planes, sizeof(double) * DEG * (size_t)n_streams, cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

int mantis_markov_update_errors(mantis_ctx* c, const double* planes, int n_streams, const int* stream_of, const double* yaw,
                                double* err, int n_hyps, int mem)
{
    if (!c || !planes || !stream_of || !yaw || !err || n_streams <= 0 || n_hyps < 0 ||
        (mem != MANTIS_MEM_HOST && mem != MANTIS_MEM_DEVICE))
        return MANTIS_ERR_INVALID;
    if (n_hyps == 0) return MANTIS_OK;
    MCHECK(c, cudaSetDevice(c->device));
    if (mem == MANTIS_MEM_HOST)
        for (int i = 0; i < n_hyps; ++i)
            if (stream_of[i] < 0 || stream_of[i] >= n_streams) return MANTIS_ERR_INVALID;
    const double *d_planes, *d_yaw;
    const int* d_so;
    double* d_err = err;
    int rc;
    if ((rc = to_device(c, c->leg_c, planes, (size_t)DEG * n_streams, mem, &d_planes))) return rc;
    if ((rc = to_device(c, c->leg_a, yaw, (size_t)n_hyps, mem, &d_yaw))) return rc;
    if ((rc = to_device(c, c->leg_b, stream_of, (size_t)n_hyps, mem, &d_so))) return rc;
    if (mem == MANTIS_MEM_HOST) {
        if ((rc = ctx_reserve(c, c->stage_out1, sizeof(double) * (size_t)n_hyps))) return rc;
        d_err = (double*)c->stage_out1.p;
        MCHECK(c, cudaMemcpyAsync(d_err, err, sizeof(double) * (size_t)n_hyps, cudaMemcpyHostToDevice, c->stream));
    }
    k_markov_errors<<<(n_hyps + 255) / 256, 256, 0, c->stream>>>(d_planes, d_so, d_yaw, d_err, n_hyps);
    c->launches++;
    MCHECK(c, cudaGetLastError());
    if (mem == MANTIS_MEM_HOST) MCHECK(c, cudaMemcpyAsync(err, d_err, sizeof(double) * (size_t)n_hyps, cudaMemcpyDeviceToHost, c->stream));
    MCHECK(c, cudaStreamSynchronize(c->stream));
    return MANTIS_OK;
}

}  // extern "C"
