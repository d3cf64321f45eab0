// fastmath.cuh -- fp64 reciprocal and reciprocal square root that stay on the DFMA pipe.
#pragma once

// fp64 building blocks that stay on the DFMA pipe.
// The projection is issue-bound on fp64, and most of its instructions were the library division,
// sqrt and atan.  These replacements seed from the fp32 special-function unit and finish with
// Newton steps in fp64 (relative error a few 1e-16, i.e. the same last-bit neighbourhood as the
// library routines).  search.cu's atan adds a 128-entry table of (theta_i, tan theta_i) and the addition
// formula, so only a degree-7 odd polynomial on |x| <= 0.0123 remains.
__device__ __forceinline__ double rcp_seed(double x)  // MUFU.RCP64H: relative error 2^-23
{
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}
__device__ __forceinline__ double rsqrt_seed(double x)  // MUFU.RSQ64H: relative error 2^-22
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}
__device__ __forceinline__ double fast_rcp(double z)  // z normal and finite
{
    double y = rcp_seed(z);
    double e = fma(-z, y, 1.0);
    y = fma(y, e, y);
    e = fma(-z, y, 1.0);
    return fma(y, e, y);
}
__device__ __forceinline__ double fast_rsqrt(double x)  // x normal and finite
{
    double r = rsqrt_seed(x);
    double e = fma(-(x * r), r, 1.0);
    r = fma(0.5 * r, e, r);
    e = fma(-(x * r), r, 1.0);
    return fma(0.5 * r, e, r);
}
