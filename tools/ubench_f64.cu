// fp64 / conversion / special-function throughput on the pipes k_search leans on (tuning aid).
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2048
template <int OP> __global__ void k(double* out, double seed)
{
    double a = threadIdx.x * 1e-3 + seed, b = a * 1.5 + 1, c = a + 0.25, d = a + 7;
    const double e = 1.0000001, f = 1e-9;
    float fa = (float)a, fb = (float)b, fc = (float)c, fd = (float)d;
    int ia = 0, ib = 0, ic = 0, id = 0;
    #pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
        #pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (OP == 0) { a = fma(a, e, f); b = fma(b, e, f); c = fma(c, e, f); d = fma(d, e, f); }
            if (OP == 1) { a = a * e; b = b * e; c = c * e; d = d * e; }
            if (OP == 2) { a = a + f; b = b + f; c = c + f; d = d + f; }
            if (OP == 3) { asm("rcp.approx.ftz.f64 %0, %0;" : "+d"(a)); asm("rcp.approx.ftz.f64 %0, %0;" : "+d"(b));
                           asm("rcp.approx.ftz.f64 %0, %0;" : "+d"(c)); asm("rcp.approx.ftz.f64 %0, %0;" : "+d"(d)); }
            if (OP == 4) { ia += __double2int_rn(a); ib += __double2int_rn(b); ic += __double2int_rn(c); id += __double2int_rn(d);
                           a += 1.0; b += 1.0; c += 1.0; d += 1.0; }
            if (OP == 5) { fa += (float)a; fb += (float)b; fc += (float)c; fd += (float)d; a += 1.0; b += 1.0; c += 1.0; d += 1.0; }
            if (OP == 6) { fa = fmaf(fa, 1.0001f, 1e-5f); fb = fmaf(fb, 1.0001f, 1e-5f); fc = fmaf(fc, 1.0001f, 1e-5f); fd = fmaf(fd, 1.0001f, 1e-5f); }
            if (OP == 7) { ia += (a > e) ? 1 : 0; ib += (b > e) ? 1 : 0; ic += (c > e) ? 1 : 0; id += (d > e) ? 1 : 0;
                           a += 1.0; b += 1.0; c += 1.0; d += 1.0; }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a + b + c + d + fa + fb + fc + fd + ia + ib + ic + id;
}
template <int OP> void run(const char* name, double* d, double per_iter)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int blocks = 148 * 8, threads = 256;
    k<OP><<<blocks, threads>>>(d, 1);
    cudaEventRecord(e0);
    k<OP><<<blocks, threads>>>(d, 2);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)blocks * threads * ITERS * 16 * per_iter;
    printf("%-28s %8.3f ms  %6.1f thread-ops/clk/SM (at 1.965 GHz)\n", name, ms, ops / (ms * 1e-3) / 148 / 1.965e9);
}
int main()
{
    double* d; cudaMalloc(&d, 148 * 8 * 256 * 8);
    run<0>("dfma", d, 1); run<1>("dmul", d, 1); run<2>("dadd", d, 1); run<3>("mufu.rcp64h", d, 1);
    run<4>("f2i.f64 + dadd (pairs)", d, 1); run<5>("f2f.f32.f64 + fadd + dadd", d, 1); run<6>("ffma", d, 1);
    run<7>("dsetp + sel + dadd (pairs)", d, 1);
    return cudaDeviceSynchronize() != cudaSuccess;
}
