// no_alloc.cu -- vlut_mul_mat_no_alloc: the whole op (quantise + ternary GEMM + fp32 scale epilogue) for callers that cannot
// allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// It is a variant of the SIMPLE anchor (gemm_simple.cu) with its own quantise step.  The kernels' scratch -- the quantised
// activations, which every other entry point keeps in the context workspace that vlut_ctx_create / vlut_quantize_act size
// with cudaMalloc -- lives in ONE file-scope __device__ array of VLUT_NO_ALLOC_MAX_SCRATCH_BYTES (4 MiB).  Every GPU that
// loads the library holds that array until the process exits.  The entry point needs no vlut_ctx, calls no cudaMalloc /
// cudaFree / allocating library routine, and rejects shapes whose scratch does not fit.
//
//   scratch layout:   scales f32 [T] (rounded up to 256 B) | q int8 [T][Kpad], Kpad = round_up(K, 128), zero padded
//
// Every call shares the array and finds the previous call's contents in it: both kernels write every byte they later read
// (the K padding included), nothing has to start at zero.  Two calls that overlap on the GPU (different streams, or
// different host threads on different streams) RACE on the scratch; calls on one stream are ordered by the stream.
//
// Arithmetic: identical to vlut_mul_mat (bit-exact): quant_math.cuh for the activations
// (quantize_row_i8_v_tile, ggml-cpu-quants-vlut.c:952-973), int32 sums per K-group wrapped to int16 at every group boundary
// (ggml-cpu-quants-vlut.c:807-826, :80-87), (float) acc * scale[t] (* w_scale) (ggml-cpu-quants-vlut.c:1252-1261).
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "quant_math.cuh"
#include "vlut_common.cuh"

namespace vlut {
namespace {

constexpr size_t NA_BYTES = VLUT_NO_ALLOC_MAX_SCRATCH_BYTES;
// In host code this name is only a handle for cudaGetSymbolAddress -- NOT a device address.
static __device__ __align__(256) unsigned char g_na_scratch[NA_BYTES];

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mutex;
unsigned char *g_na_ptr[NA_MAX_DEVICES];   // device address of g_na_scratch on every device seen so far

constexpr int NQ = 256;

// one CTA per token: signed row maximum -> scale, then q = clamp(round(x / scale)); the K padding is written as zeros
__global__ void __launch_bounds__(NQ) na_quantize_kernel(const float *__restrict__ x, int64_t ldx, int64_t K, int64_t Kpad,
                                                         float *__restrict__ scales, int8_t *__restrict__ q) {
    const int64_t t = blockIdx.x;
    const int tid = threadIdx.x;
    const float *xr = x + t * ldx;
    __shared__ float s_red[NQ / 32];
    __shared__ double s_is;
    float mx = -INFINITY;
    for (int64_t k = tid; k < K; k += NQ) mx = fmaxf(mx, xr[k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((tid & 31) == 0) s_red[tid >> 5] = mx;
    __syncthreads();
    if (tid == 0) {
        float m = s_red[0];
        for (int i = 1; i < NQ / 32; i++) m = fmaxf(m, s_red[i]);
        double is;
        float sc;
        quant_scale(m, is, sc);
        s_is = is;
        scales[t] = sc;
    }
    __syncthreads();
    const double is = s_is;
    int8_t *qr = q + t * Kpad;
    for (int64_t k = tid; k < Kpad; k += NQ) qr[k] = (k < K) ? (int8_t) quant1_exact(xr[k], is) : (int8_t) 0;
}

// one thread per output element (threadIdx.x walks n: coalesced stores), as the SIMPLE anchor
__global__ void __launch_bounds__(128) na_gemm_kernel(const uint8_t *__restrict__ w, const float *__restrict__ scales,
                                                      const int8_t *__restrict__ q, int64_t K, int64_t Kpad, int64_t N,
                                                      int gsize, float w_scale, float *__restrict__ out, int64_t ldo) {
    const int64_t n = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t t = blockIdx.y;
    if (n >= N) return;
    const int8_t *__restrict__ qr = q + t * Kpad;
    int32_t total = 0;
    for (int64_t k0 = 0; k0 < K; k0 += gsize) {
        const int64_t k1 = (k0 + gsize < K) ? k0 + gsize : K;
        int32_t s = 0;
        for (int64_t k = k0; k < k1; k += 4) {   // K % 4 == 0 for every vlut type
            const uint32_t word = *reinterpret_cast<const uint32_t *>(w + t2_word_offset(n, k, Kpad));
            const uint32_t qq = *reinterpret_cast<const uint32_t *>(qr + k);
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int c = (int) ((word >> t2_shift(k + j)) & 3u) - 1;
                s += c * (int) (int8_t) (qq >> (8 * j));
            }
        }
        total += wrap16(s);
    }
    float v = (float) total * scales[t];
    if (w_scale != 1.0f) v *= w_scale;
    out[t * ldo + n] = v;
}

int na_fail(char *err, size_t err_len, int code, const char *fmt, ...) {
    if (err && err_len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace
}  // namespace vlut

using namespace vlut;

extern "C" int vlut_mul_mat_no_alloc(int device, void *stream, int type, int64_t K, int64_t N, int64_t T,
                                     const void *t2_dev, const float *x_dev, int64_t ldx, float *out_dev, int64_t ldo,
                                     float w_scale, char *err, size_t err_len) {
    if (err && err_len) err[0] = 0;
    if (!t2_dev || !x_dev || !out_dev)
        return na_fail(err, err_len, VLUT_ERR_INVALID, "vlut_mul_mat_no_alloc: null device pointer");
    if (T <= 0 || ldx < K || ldo < N || !vlut_op_supported(type, K, N, T))
        return na_fail(err, err_len, VLUT_ERR_INVALID,
                       "vlut_mul_mat_no_alloc: bad type / shape (type %d, K %lld, N %lld, T %lld, ldx %lld, ldo %lld)", type,
                       (long long) K, (long long) N, (long long) T, (long long) ldx, (long long) ldo);
    if (device < 0 || device >= NA_MAX_DEVICES)
        return na_fail(err, err_len, VLUT_ERR_INVALID, "vlut_mul_mat_no_alloc: device ordinal %d out of range", device);
    if (T > 65535)
        return na_fail(err, err_len, VLUT_ERR_UNSUPPORTED, "vlut_mul_mat_no_alloc: T = %lld tokens exceeds 65535", (long long) T);
    const int64_t Kpad = round_up(K, BLOCK_K);
    const size_t scales_bytes = (size_t) round_up(T * 4, 256);
    const size_t need = scales_bytes + (size_t) T * (size_t) Kpad;
    if (need > NA_BYTES)
        return na_fail(err, err_len, VLUT_ERR_UNSUPPORTED,
                       "vlut_mul_mat_no_alloc: T = %lld tokens x K = %lld needs %zu bytes of scratch, the static scratch array "
                       "holds VLUT_NO_ALLOC_MAX_SCRATCH_BYTES = %zu (T * (round_up(K, 128) + 4) must fit); split the tokens "
                       "or use vlut_mul_mat", (long long) T, (long long) K, need, NA_BYTES);

    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
    if (e != cudaSuccess)
        return na_fail(err, err_len, VLUT_ERR_CUDA, "vlut_mul_mat_no_alloc: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    int rc = VLUT_OK;
    unsigned char *base = nullptr;
    {   // set-up, once per device, with that device current: the array's device address
        std::lock_guard<std::mutex> lock(g_na_mutex);
        if (!g_na_ptr[device]) {
            void *ptr = nullptr;
            e = cudaGetSymbolAddress(&ptr, g_na_scratch);
            if (e != cudaSuccess || !ptr)
                rc = na_fail(err, err_len, VLUT_ERR_CUDA, "vlut_mul_mat_no_alloc: cudaGetSymbolAddress on device %d: %s", device,
                             cudaGetErrorString(e));
            else
                g_na_ptr[device] = (unsigned char *) ptr;
        }
        base = g_na_ptr[device];
    }
    if (rc == VLUT_OK) {
        float *scales = reinterpret_cast<float *>(base);
        int8_t *q = reinterpret_cast<int8_t *>(base + scales_bytes);
        cudaStream_t s = (cudaStream_t) stream;
        na_quantize_kernel<<<(unsigned) T, NQ, 0, s>>>(x_dev, ldx, K, Kpad, scales, q);
        dim3 grid((unsigned) ceil_div(N, 128), (unsigned) T);
        na_gemm_kernel<<<grid, 128, 0, s>>>((const uint8_t *) t2_dev, scales, q, K, Kpad, N, group_size(type), w_scale, out_dev,
                                            ldo);
        e = cudaGetLastError();
        if (e != cudaSuccess)
            rc = na_fail(err, err_len, VLUT_ERR_CUDA, "vlut_mul_mat_no_alloc: kernel launch: %s", cudaGetErrorString(e));
    }
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    return rc;
}
