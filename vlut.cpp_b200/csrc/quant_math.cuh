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

// Four elements -> four packed int8, ~8 instructions per element.  Round-to-nearest-even through the 1.5 * 2^23 trick (the
// low mantissa byte of pc + MAGIC is the two's complement int8), which differs from round-half-away only AT a tie; the
// same closeness test as quant1 (distance to a k + 0.5 boundary <= 1e-4, i.e. |pc - rounded| >= 0.4999) sends ties and
// near-ties to the exact double path, for all four elements at once (about 1 call in 1250).  The clamp to [-128, 127] is
// applied before rounding: it commutes with rounding because its bounds are integers.
__device__ __forceinline__ uint32_t quant4(float4 v, double is, float isf) {
    constexpr float MAGIC = 12582912.0f;
    const float p0 = fminf(fmaxf(v.x * isf, -128.0f), 127.0f), p1 = fminf(fmaxf(v.y * isf, -128.0f), 127.0f);
    const float p2 = fminf(fmaxf(v.z * isf, -128.0f), 127.0f), p3 = fminf(fmaxf(v.w * isf, -128.0f), 127.0f);
    const float y0 = p0 + MAGIC, y1 = p1 + MAGIC, y2 = p2 + MAGIC, y3 = p3 + MAGIC;
    // |p - round(p)|: both subtractions are exact
    const float d = fmaxf(fmaxf(fabsf(p0 - (y0 - MAGIC)), fabsf(p1 - (y1 - MAGIC))),
                          fmaxf(fabsf(p2 - (y2 - MAGIC)), fabsf(p3 - (y3 - MAGIC))));
    if (d >= 0.4999f) {
        const int q0 = quant1_exact(v.x, is), q1 = quant1_exact(v.y, is), q2 = quant1_exact(v.z, is), q3 = quant1_exact(v.w, is);
        return (uint32_t) (q0 & 0xff) | ((uint32_t) (q1 & 0xff) << 8) | ((uint32_t) (q2 & 0xff) << 16) | ((uint32_t) (q3 & 0xff) << 24);
    }
    const uint32_t lo = __byte_perm(__float_as_uint(y0), __float_as_uint(y1), 0x0040);
    const uint32_t hi = __byte_perm(__float_as_uint(y2), __float_as_uint(y3), 0x0040);
    return __byte_perm(lo, hi, 0x5410);
}

// sum of the four int8 of a packed word, and of their absolute values
__device__ __forceinline__ int sum4(uint32_t pk) { return __dp4a((int) pk, 0x01010101, 0); }
__device__ __forceinline__ int abssum4(uint32_t pk) {
    // per byte: +1 (0x01) for q >= 0, -1 (0xff) for q < 0
    const uint32_t sg = ((pk >> 7) & 0x01010101u) * 0xfeu + 0x01010101u;
    return __dp4a((int) pk, (int) sg, 0);
}

}  // namespace vlut
