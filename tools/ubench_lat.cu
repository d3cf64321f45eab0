// Dependent-issue latency of the fp64 / conversion / special-function instructions k_search chains (tuning aid).
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
template <int OP> __global__ void k(double* out, double seed, long long* cyc)
{
    double a = threadIdx.x * 1e-3 + seed;
    const double e = 1.0000001, f = 1e-9;
    float fa = (float)a;
    int ia = threadIdx.x;
    long long t0 = clock64();
    #pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
        #pragma unroll
        for (int u = 0; u < 16; ++u) {
            if (OP == 0) a = fma(a, e, f);
            if (OP == 1) a = a * e;
            if (OP == 2) asm("rcp.approx.ftz.f64 %0, %0;" : "+d"(a));
            if (OP == 3) { ia = __double2int_rn(a); a = __hiloint2double(0x40000000 | (ia & 0xff), ia); }
            if (OP == 4) { fa = (float)a; a = __hiloint2double(__float_as_int(fa), 0); }
            if (OP == 5) fa = fmaf(fa, 1.0001f, 1e-5f);
            if (OP == 6) { ia = (a > e) ? ia + 1 : ia; a = __hiloint2double(0x40000000 | (ia & 0xff), ia); }
            if (OP == 7) ia = ia * 3 + 1;
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a + fa + ia;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int OP> void run(const char* name, double* d, long long* c)
{
    k<OP><<<1, 32>>>(d, 1, c);
    k<OP><<<1, 32>>>(d, 2, c);
    long long h; cudaMemcpy(&h, c, 8, cudaMemcpyDeviceToHost);
    printf("%-34s %6.2f cycles per dependent op\n", name, (double)h / (ITERS * 16.0));
}
int main()
{
    double* d; cudaMalloc(&d, 4096 * 8); long long* c; cudaMalloc(&c, 8);
    run<0>("dfma", d, c); run<1>("dmul", d, c); run<2>("mufu.rcp64h", d, c); run<3>("f2i.f64 (+mov pair)", d, c);
    run<4>("f2f.f32.f64 (+mov)", d, c); run<5>("ffma", d, c); run<6>("dsetp + sel (+mov)", d, c); run<7>("imad", d, c);
    return cudaDeviceSynchronize() != cudaSuccess;
}
