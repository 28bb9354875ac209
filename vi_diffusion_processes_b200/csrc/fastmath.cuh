// float64 reciprocal / square root / division without the IEEE slow paths.
//
// nvcc expands `a / b` and `sqrt(x)` on doubles into ~30-instruction sequences with a guarded slow path
// (denormals, infinities) around the same hardware seed used here.  The sweeps of this library are bound by
// instruction issue, not by the FP64 pipe, and every operand is a variance, a precision or a Cholesky pivot
// (normal range, positive), so the guards buy nothing.  These versions: MUFU seed (rcp/rsqrt.approx.ftz.f64,
// >= 20 good bits) + two Newton / Goldschmidt steps (2^-20 -> 2^-40 -> 2^-80) + one residual correction.
// Error <= 1 ulp (measured by tests/test_gpu_fastmath.py through vdp_test_fastmath); zero, negative,
// denormal or non-finite operands give NaN/Inf garbage, which the callers' positive-definiteness checks
// report through the status word before the value is used.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace vdp {

#define VDP_FM __host__ __device__ __forceinline__

// 1 / x
VDP_FM double rcp_fast(double x) {
#ifdef __CUDA_ARCH__
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    return r;
#else
    return 1.0 / x;
#endif
}

// a / b given r ~ 1 / b  (one residual correction: the result is the correctly rounded quotient except in
// rare half-way cases, where it is its neighbour)
VDP_FM double div_r(double a, double b, double r) {
#ifdef __CUDA_ARCH__
    const double q = a * r;
    return fma(fma(-b, q, a), r, q);
#else
    (void)r;
    return a / b;
#endif
}

VDP_FM double div_fast(double a, double b) { return div_r(a, b, rcp_fast(b)); }

// s = sqrt(x) and rs = 1 / sqrt(x) from one seed (coupled Goldschmidt iteration)
VDP_FM void sqrt_rsqrt_fast(double x, double& s, double& rs) {
#ifdef __CUDA_ARCH__
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double g = x * r, h = 0.5 * r;
    double e = fma(-g, h, 0.5);
    g = fma(g, e, g);
    h = fma(h, e, h);
    e = fma(-g, h, 0.5);
    g = fma(g, e, g);
    h = fma(h, e, h);
    g = fma(fma(-g, g, x), h, g);     // g + (x - g^2) / (2 g)
    e = fma(-g, h, 0.5);
    h = fma(h, e, h);
    s = g;
    rs = h + h;
#else
    s = sqrt(x);
    rs = 1.0 / s;
#endif
}

VDP_FM double sqrt_fast(double x) {
#ifdef __CUDA_ARCH__
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double g = x * r, h = 0.5 * r;
    double e = fma(-g, h, 0.5);
    g = fma(g, e, g);
    h = fma(h, e, h);
    e = fma(-g, h, 0.5);
    g = fma(g, e, g);
    h = fma(h, e, h);
    return fma(fma(-g, g, x), h, g);
#else
    return sqrt(x);
#endif
}

}  // namespace vdp
