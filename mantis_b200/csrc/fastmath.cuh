// fastmath.cuh -- fp64 reciprocal and reciprocal square root that stay on the DFMA pipe.
#pragma once

// fp64 building blocks that stay on the DFMA pipe.
// The projection is issue-bound on fp64, and most of its instructions were the library division,
// sqrt and atan.  These replacements seed from the fp32 special-function unit and finish with
// Newton steps in fp64 (relative error a few 1e-16, i.e. the same last-bit neighbourhood as the
// library routines).  search.cu's atan adds a 128-entry table of (theta_i, tan theta_i) and the addition
// formula, so only a degree-7 odd polynomial on |x| <= 0.0123 remains.
#ifdef MANTIS_HOST_EMUL  // tools/rpp_emul only (a host build of the pose code for diagnostics); never in the library
__device__ __forceinline__ double rcp_seed(double x) { return (double)(1.0f / (float)x); }
__device__ __forceinline__ double rsqrt_seed(double x) { return (double)(float)(1.0 / sqrt(x)); }
#else
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
#endif
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

// Cubic-step variants (one instruction fewer / three fewer), used by the projection of search.cu.
__device__ __forceinline__ double fast_rcp3(double z)  // z normal and finite
{
    // y (1 + e + e^2) with e = 1 - z y; e ~ 2^-23, so the result is good to e^3 ~ 2^-69
    const double y = rcp_seed(z);
    const double e = fma(-z, y, 1.0);
    return fma(y, fma(e, e, e), y);
}
__device__ __forceinline__ double fast_rsqrt3(double x)  // x normal and finite
{
    // r (1 + e/2 + 3 e^2 / 8) with e = 1 - x r^2; e ~ 2^-21, so the result is good to ~ 2^-64
    const double r = rsqrt_seed(x);
    const double e = fma(-(x * r), r, 1.0);
    return fma(r, fma(e, 0.375, 0.5) * e, r);
}
