// gemm_simple.cu -- VLUT_ALGO_SIMPLE: the in-library correctness anchor.
//
// One thread per output element, scalar decode of the T2 words, int32 accumulation per K-group, int16 wrap at
// every group boundary, fp32 epilogue.  Deliberately naive (no shared memory, no tensor cores): it exists so
// that the fast kernels can be cross-checked on the device at sizes the CPU oracle cannot reach, and it is
// the fallback for shapes whose K-groups do not align with the 128-wide T2 blocks.
//   acc32[t][n] = sum_g wrap16( sum_{k in g} w[n][k] * q[t][k] )      (ggml-cpu-quants-vlut.c:807-826, :80-87)
//   out[t][n]   = (float) acc32 * scale[t]                            (ggml-cpu-quants-vlut.c:1252-1261)
#include "vlut_common.cuh"

namespace vlut {

__global__ void __launch_bounds__(128) gemm_simple_kernel(GemmArgs p) {
    // threadIdx.x walks n (coalesced output), blockIdx.y walks t
    const int64_t n = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t t = blockIdx.y;
    if (n >= p.N || t >= p.T) return;
    const int64_t Kpad = p.act.Kpad;
    const int8_t *__restrict__ qr = p.act.q_rm + t * Kpad;
    const int gsize = p.act.gsize;
    int32_t total = 0;
    for (int64_t k0 = 0; k0 < p.K; k0 += gsize) {
        const int64_t k1 = (k0 + gsize < p.K) ? k0 + gsize : p.K;
        int32_t s = 0;
        for (int64_t k = k0; k < k1; k += 4) {  // K % 4 == 0 for every vlut type; 4 k share byte lane pattern
            const uint32_t word = *reinterpret_cast<const uint32_t *>(p.w + t2_word_offset(n, k, Kpad));
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int c = (int) ((word >> t2_shift(k + j)) & 3u) - 1;
                s += c * (int) qr[k + j];
            }
        }
        total += wrap16(s);
    }
    float v = (float) total * p.act.scales[t];
    if (p.w_scale != 1.0f) v *= p.w_scale;
    p.out[t * p.ldo + n] = v;
}

cudaError_t launch_gemm_simple(const GemmArgs &g, cudaStream_t s) {
    dim3 grid((unsigned) ceil_div(g.N, 128), (unsigned) g.T);
    gemm_simple_kernel<<<grid, 128, 0, s>>>(g);
    return cudaGetLastError();
}

// out[t][n] = x[t][n] * s: GGML_OP_MUL by a one-element f32 tensor (the `.scale` of a ternary weight,
// src/models/bitnet.cpp:33-34; ggml_compute_forward_mul, ggml-cpu/binary-ops.cpp:152) when it could not be folded
// into the GEMM epilogue.  One fp32 rounding, as on the CPU.
__global__ void __launch_bounds__(256) scale_rows_kernel(const float *x, int64_t ldx, float *out,
                                                         int64_t ldo, int64_t N, int64_t T, float s) {
    const int64_t t = blockIdx.y;
    for (int64_t n = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; n < N; n += (int64_t) gridDim.x * blockDim.x)
        out[t * ldo + n] = x[t * ldx + n] * s;
}

cudaError_t launch_scale_rows(const float *x, int64_t ldx, float *out, int64_t ldo, int64_t N, int64_t T, float sc,
                              cudaStream_t s) {
    for (int64_t t0 = 0; t0 < T; t0 += 65535) {
        const int64_t tn = T - t0 < 65535 ? T - t0 : 65535;
        dim3 grid((unsigned) (ceil_div(N, 256) < 64 ? ceil_div(N, 256) : 64), (unsigned) tn);
        scale_rows_kernel<<<grid, 256, 0, s>>>(x + t0 * ldx, ldx, out + t0 * ldo, ldo, N, tn, sc);
    }
    return cudaGetLastError();
}

}  // namespace vlut
