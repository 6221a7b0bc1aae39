// quant_math.cuh -- the per-element arithmetic of quantize_row_i8_v_tile (reference
// ggml/src/ggml-cpu/ggml-cpu-quants-vlut.c:952-973), shared by the stand-alone quantise kernel (quantize.cu) and the
// kernels that quantise their own activations (gemm_gemv.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vlut {

// row maximum (signed, no fabs) -> inverse scale `is` (double) and the fp32 scale the epilogue multiplies by
__device__ __forceinline__ void quant_scale(float row_max, double &is, float &scale) {
    double dm = (double) row_max;
    if (!(dm > 1e-5)) dm = 1e-5;  // MAX(eps, .) with eps = 1e-5
    const double sd = dm / 127;
    is = 1e0 / (sd > 1e-5 ? sd : 1e-5);
    scale = (float) sd;
}

__device__ __forceinline__ int quant1_exact(float x, double is) {
    // clamp(round((double) x * is), -128, 127) with round = half away from zero, exactly: the clamp is applied first (it
    // commutes with rounding at integers), then truncation and an exact fractional test
    double p = (double) x * is;
    p = fmin(fmax(p, -129.0), 128.0);
    int ti = __double2int_rz(p);
    const double fr = p - (double) ti;
    ti += (fr >= 0.5) - (fr <= -0.5);
    return max(-128, min(127, ti));
}

// Same value, fp32 first: pf = x * (float) is differs from the exact double product by at most |p| * 2^-23 <= 1.6e-5 inside
// the clamp range, so whenever pf is further than 1e-4 from a rounding boundary (k + 0.5) both round the same way; the
// few elements closer than that (about 1 in 5000) take the double path.  The FP64 pipe (conversions included) was ~2 us of
// the kernel at T = 512.
__device__ __forceinline__ int quant1(float x, double is, float isf) {
    float pf = fminf(fmaxf(x * isf, -129.0f), 128.0f);
    const float tr = truncf(pf);
    const float fr = fabsf(pf - tr);   // exact
    if (fabsf(fr - 0.5f) <= 1e-4f) return quant1_exact(x, is);
    const int ti = (int) tr + ((fr > 0.5f) ? (pf < 0.0f ? -1 : 1) : 0);
    return max(-128, min(127, ti));
}

}  // namespace vlut
