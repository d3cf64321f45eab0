// Instruction-throughput microbenchmark for the integer ops the image kernels lean on (tuning aid).
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#define ITERS 4096
template <int OP> __global__ void k(unsigned* out, unsigned seed)
{
    unsigned a = threadIdx.x + seed, b = a * 3 + 1, c = a ^ 0x55, d = a + 7;
    unsigned e = b + 11, f = c + 13, g = d + 17, h = a + 19;
    #pragma unroll 1
    for (int i = 0; i < ITERS; ++i) {
        #pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (OP == 0) { a = a * 84u + e; b = b * 84u + f; c = c * 84u + g; d = d * 84u + h; }
            if (OP == 1) { a = __dp4a(a, 0x00151615u, e); b = __dp4a(b, 0x00151615u, f); c = __dp4a(c, 0x00151615u, g); d = __dp4a(d, 0x00151615u, h); }
            if (OP == 2) { a = __dp2a_lo(a, 0x1615u, e); b = __dp2a_lo(b, 0x1615u, f); c = __dp2a_lo(c, 0x1615u, g); d = __dp2a_lo(d, 0x1615u, h); }
            if (OP == 3) { a = __byte_perm(a, e, 0x5140); b = __byte_perm(b, f, 0x5140); c = __byte_perm(c, g, 0x5140); d = __byte_perm(d, h, 0x5140); }
            if (OP == 4) { a = (a & e) ^ 0x00ff00ffu; b = (b & f) ^ 0x00ff00ffu; c = (c & g) ^ 0x00ff00ffu; d = (d & h) ^ 0x00ff00ffu; }
            if (OP == 5) { a = __funnelshift_r(a, e, 8); b = __funnelshift_r(b, f, 8); c = __funnelshift_r(c, g, 8); d = __funnelshift_r(d, h, 8); }
            if (OP == 6) { a = a + e + 0x01000100u; b = b + f + 0x01000100u; c = c + g + 0x01000100u; d = d + h + 0x01000100u; }
            if (OP == 7) { a = __sad((int)a, (int)e, b); b = __sad((int)b, (int)f, c); c = __sad((int)c, (int)g, d); d = __sad((int)d, (int)h, a); }
            if (OP == 8) { a = __vabsdiffu2(a, e); b = __vabsdiffu2(b, f); c = __vabsdiffu2(c, g); d = __vabsdiffu2(d, h); }
            if (OP == 9) { __half2 x = *(__half2*)&a, y = *(__half2*)&e; x = __hfma2(x, y, y); a = *(unsigned*)&x;
                           __half2 x2 = *(__half2*)&b, y2 = *(__half2*)&f; x2 = __hfma2(x2, y2, y2); b = *(unsigned*)&x2;
                           __half2 x3 = *(__half2*)&c, y3 = *(__half2*)&g; x3 = __hfma2(x3, y3, y3); c = *(unsigned*)&x3;
                           __half2 x4 = *(__half2*)&d, y4 = *(__half2*)&h; x4 = __hfma2(x4, y4, y4); d = *(unsigned*)&x4; }
            if (OP == 10) { a = __vcmpgtu2(a, e); b = __vcmpgtu2(b, f); c = __vcmpgtu2(c, g); d = __vcmpgtu2(d, h); }
            if (OP == 11) { a = max(a, e) ; b = max(b, f); c = max(c, g); d = max(d, h); a ^= 1; b ^= 1; c ^= 1; d ^= 1; }
            if (OP == 12) { a = a * 84u + e; b = (b & f) ^ 0x00ff00ffu; c = c * 84u + g; d = (d & h) ^ 0x00ff00ffu; }
            if (OP == 13) { a = __vadd2(a, e); b = __vsub2(b, f); c = __vadd2(c, g); d = __vsub2(d, h); }
            if (OP == 14) { a = __popc(a) + e; b = __popc(b) + f; c = __popc(c) + g; d = __popc(d) + h; }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a + b + c + d;
}
template <int OP> void run(const char* name, unsigned* d)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int blocks = 148 * 8, threads = 256;
    k<OP><<<blocks, threads>>>(d, 1);
    cudaEventRecord(e0);
    k<OP><<<blocks, threads>>>(d, 2);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)blocks * threads * ITERS * 16;
    printf("%-14s %8.3f ms  %7.2f Gop/s/SM-clk-equiv: %6.1f thread-ops/clk/SM (at 1.965 GHz)\n", name, ms, ops / ms / 1e6,
           ops / (ms * 1e-3) / 148 / 1.965e9);
}
int main()
{
    unsigned* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("imad", d); run<1>("dp4a", d); run<2>("dp2a", d); run<3>("prmt", d); run<4>("lop3", d); run<5>("shf", d);
    run<6>("iadd3", d); run<7>("sad", d); run<8>("vabsdiffu2", d); run<9>("hfma2", d); run<10>("vcmpgtu2", d);
    run<11>("max+xor", d); run<12>("imad+lop mix", d); run<13>("vadd2/vsub2", d); run<14>("popc+add", d);
    return cudaDeviceSynchronize() != cudaSuccess;
}
