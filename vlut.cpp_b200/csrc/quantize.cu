// quantize.cu -- stage 1: per-token int8 activation quantisation.
//
// Bit-exact restatement, for sm_100a, of quantize_row_i8_v_tile (reference
// ggml/src/ggml-cpu/ggml-cpu-quants-vlut.c:952-973):
//     m  = max(1e-5, max_k x[k])            SIGNED max -- no fabs
//     s  = m / 127                          (double)
//     is = 1 / max(s, 1e-5)                 (double)
//     q  = clamp(round((double)x * is), -128, 127)     round = half away from zero
//     scale = (float) s
// FP64 throughput is irrelevant here (T*K elements); the kernel is a pure HBM/L2 stream: one CTA per token,
// 128-bit loads of x, 32-bit stores of q in two layouts (row-major, and pre-swizzled K-major UMMA operand tiles
// that a TMA bulk copy drops into shared memory ready for tcgen05.mma), plus per (token, K-group): the sum of q
// that undoes the +1 offset of the u8 weight codes, and whether sum|q| rules out any int16 wrap.
#include "ptx.cuh"
#include "quant_math.cuh"
#include "vlut_common.cuh"

namespace vlut {

constexpr int QT = 256;

__global__ void __launch_bounds__(QT) quantize_kernel(const float *__restrict__ x, int64_t ldx, int64_t K, int64_t T,
                                                      ActView a, int want_rm) {
    const int t = blockIdx.x;
    const int tid = threadIdx.x;
    // programmatic dependent launch: let the GEMM that follows start its prologue and its weight prefetch now; it
    // orders itself against this kernel's results with griddepcontrol.wait
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __shared__ float s_red[QT / 32];
    __shared__ int s_gsum[64];
    __shared__ int s_gabs[64];
    __shared__ double s_is;

    const int64_t Kpad = a.Kpad;
    const bool live = t < T;
    const float *xr = x + (int64_t) t * ldx;

    for (int i = tid; i < 64; i += QT) { s_gsum[i] = 0; s_gabs[i] = 0; }

    // pass 1: signed max
    float mx = -INFINITY;
    if (live) {
        const bool vec = ((reinterpret_cast<uintptr_t>(xr) & 15) == 0);
        if (vec) {
            for (int64_t k = (int64_t) tid * 4; k + 3 < K; k += QT * 4) {
                const float4 v = *reinterpret_cast<const float4 *>(xr + k);
                mx = fmaxf(mx, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
            }
            for (int64_t k = (K & ~3LL) + tid; k < K; k += QT) mx = fmaxf(mx, xr[k]);
        } else {
            for (int64_t k = tid; k < K; k += QT) mx = fmaxf(mx, xr[k]);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((tid & 31) == 0) s_red[tid >> 5] = mx;
    __syncthreads();
    if (tid == 0) {
        float m = s_red[0];
        for (int i = 1; i < QT / 32; i++) m = fmaxf(m, s_red[i]);
        double dm = (double) m;
        if (!(dm > 1e-5)) dm = 1e-5;  // MAX(eps, .) with eps = 1e-5
        const double s = dm / 127;
        s_is = 1e0 / (s > 1e-5 ? s : 1e-5);
        a.scales[t] = live ? (float) s : 0.0f;
    }
    __syncthreads();
    const double is = s_is;

    // pass 2: quantise 4 consecutive k per thread-iteration
    for (int64_t k = (int64_t) tid * 4; k < Kpad; k += QT * 4) {
        int qv[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            int q = 0;
            if (live && k + j < K) {
                float v = (float) round((double) xr[k + j] * is);
                v = v > 127.f ? 127.f : (v < -128.f ? -128.f : v);
                q = (int) v;
            }
            qv[j] = q;
        }
        const uint32_t pk = (uint32_t) (qv[0] & 0xff) | ((uint32_t) (qv[1] & 0xff) << 8) |
                            ((uint32_t) (qv[2] & 0xff) << 16) | ((uint32_t) (qv[3] & 0xff) << 24);
        if (want_rm) *reinterpret_cast<uint32_t *>(a.q_rm + (int64_t) t * Kpad + k) = pk;
        // K-major UMMA tile, 128B swizzle: 16-byte chunk c of row t goes to chunk position c ^ (t % 8)
        const int64_t kb = k / BLOCK_K;
        const int kk = act_row_pos((int) (k % BLOCK_K), a.perm);   // 4 consecutive k stay 4 consecutive bytes in both orders
        const int c = kk / 16;
        const int64_t off = (((int64_t) (t / a.sw_tile) * (Kpad / BLOCK_K) + kb) * a.sw_tile + (t % a.sw_tile)) * 128 +
                            ((c ^ (t & 7)) << 4) + (kk & 15);
        *reinterpret_cast<uint32_t *>(a.q_sw + off) = pk;
        // the 32 lanes of a warp cover 128 consecutive, 128-aligned k here: one K-group (512 and 640 are multiples of
        // 128), so reduce with shuffles and touch shared memory once per warp (Kpad % 128 == 0: the loop is warp-uniform)
        int sum = qv[0] + qv[1] + qv[2] + qv[3];
        int asum = abs(qv[0]) + abs(qv[1]) + abs(qv[2]) + abs(qv[3]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sum += __shfl_xor_sync(0xffffffffu, sum, o);
            asum += __shfl_xor_sync(0xffffffffu, asum, o);
        }
        if ((tid & 31) == 0 && asum != 0) {
            atomicAdd(&s_gsum[k / a.gsize], sum);
            atomicAdd(&s_gabs[k / a.gsize], asum);
        }
    }
    __syncthreads();
    if (tid == 0) {
        int run = 0, safe = 1;
        int32_t *pre = a.gpre + (int64_t) t * (a.G + 1);
        for (int i = 0; i < a.G; i++) {
            pre[i] = run;
            run += s_gsum[i];
            if (s_gabs[i] > 32767) safe = 0;
        }
        pre[a.G] = run;
        a.nowrap[t] = safe;
        // a token tile takes the single-drain path unless one of its tokens stamps it with this call's epoch
        if (!safe) { a.tile_unsafe[t / a.NT] = a.epoch; a.tile_unsafe[a.Tpad / 16] = a.epoch; }   // [Tpad / 16]: "some tile is unsafe"
    }
}

// ---- register-resident single-pass kernel (K <= 16384, 16-byte aligned rows): the default --------------------------------
// One CTA of NTH threads per token; thread i owns the 16-byte words i, i + NTH, ... of the row (a warp instruction reads 512
// contiguous bytes), all loads issued before the first use, and the row stays in registers between the max pass and the
// quantise pass.  Output per word: one packed int8x4 into its pre-swizzled position of q_sw -- the 32 lanes of a warp fill
// exactly one 128-byte row of one swizzled tile, and (128 divides the group sizes) belong to one K-group: two redux per 128
// elements for the group sums, prefix sums by a warp scan.  Launched with the programmatic-serialisation attribute: the CTA
// is resident and past its prologue when the predecessor (the previous GEMM, which reads the workspace this kernel rewrites)
// drains; griddepcontrol.wait orders every global access after it.
//
// History (scripts/time_quant.py, profiles/r2_time_quant.txt): the first version of this kernel spent ~25 instructions per
// element (per-element round / clamp / pack) with a 64-byte stride per lane on the loads and took 6.6 us at K = 4096,
// T = 512, 20.7 us at K = 14336, T = 512, and 8.6 us for ONE token of 14336 -- it was issue-bound, not HBM-bound.  quant4
// (quant_math.cuh) is ~9 instructions per element.  A persistent variant that streamed rows through shared memory with bulk
// copies (512 threads, 3 CTA-wide barriers per token) measured SLOWER (10.1 / 23.5 us): per token the max -> scale ->
// quantise -> scan chain is ~2000 clk of latency and two CTAs per SM do not hide it; many small CTAs do.
template <int NTH, int CH>
__global__ void __launch_bounds__(NTH) quantize_reg_kernel(const float *__restrict__ x, int64_t ldx, int64_t K, int64_t T,
                                                           ActView a, int early, int want_rm) {
    const int t = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __shared__ float s_red[NTH / 32];
    __shared__ int s_gsum[64];
    __shared__ int s_gabs[64];
    if (tid < 64) { s_gsum[tid] = 0; s_gabs[tid] = 0; }
    const bool live = t < T;
    const float4 *xr = reinterpret_cast<const float4 *>(x + (int64_t) t * ldx);
    // `early` (the host knows the preceding kernel on the stream does not write x): the row is loaded, reduced and its
    // scale computed while the previous GEMM is still running; only the stores below wait for it (they overwrite the
    // workspace that GEMM reads)
    if (!early) asm volatile("griddepcontrol.wait;" ::: "memory");

    const int KW = (int) (K / 4), KPW = (int) (a.Kpad / 4);   // 16-byte words of the row / of the padded row
    float4 v[CH];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < CH; j++) {
        const int w = j * NTH + tid;
        if (live && w < KW) v[j] = __ldg(xr + w);
        else v[j] = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    }
#pragma unroll
    for (int j = 0; j < CH; j++) mx = fmaxf(mx, fmaxf(fmaxf(v[j].x, v[j].y), fmaxf(v[j].z, v[j].w)));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) s_red[warp] = mx;
    __syncthreads();
    float m = s_red[0];
#pragma unroll
    for (int i = 1; i < NTH / 32; i++) m = fmaxf(m, s_red[i]);
    double is;
    float scale;
    quant_scale(m, is, scale);
    const float isf = (float) is;
    if (early) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (tid == 0) a.scales[t] = live ? scale : 0.0f;

    const int KB = (int) (a.Kpad / BLOCK_K);
    // word `lane` of a 128-byte q_sw row -> its position: regrouping by bit-pair position (perm: word 4c + i -> 8i + c),
    // then the 128B swizzle (16-byte chunk c -> c ^ (t % 8))
    const int wp = a.perm ? ((lane & 3) * 8 + (lane >> 2)) : lane;
    const int wpos = (((wp >> 2) ^ (t & 7)) << 2) | (wp & 3);
    // Everything per word below is 32-bit and incremental (the first version of this loop spent 24 IMAD + 12 ISETP per word
    // on 64-bit addresses and a runtime division by the group size: ncu source page, profiles/r2_ncu_quant.txt).
    // Word w = j * NTH + tid sits in sub-block kb = j * (NTH / 32) + warp; sub-blocks are sw_tile * 128 bytes apart in q_sw.
    const uint32_t kb_stride = (uint32_t) a.sw_tile * 128u;
    uint8_t *dst = reinterpret_cast<uint8_t *>(a.q_sw) +
                   ((int64_t) (t / a.sw_tile) * KB * a.sw_tile + (t % a.sw_tile)) * 128 + (int64_t) warp * kb_stride + wpos * 4;
    uint8_t *dst_rm = reinterpret_cast<uint8_t *>(a.q_rm) + (int64_t) t * a.Kpad + tid * 4;
    const bool g4 = (a.gsize == 512);   // K-group = 4 sub-blocks (I2_*) or 5 (I1_*)
#pragma unroll
    for (int j = 0; j < CH; j++) {
        const int w = j * NTH + tid;
        const int kb = j * (NTH / 32) + warp;
        if (kb >= KB) break;   // warp-uniform: nothing left for this warp
        const uint32_t pk = (live && w < KW) ? quant4(v[j], is, isf) : 0u;
        *reinterpret_cast<uint32_t *>(dst + (uint32_t) (j * (NTH / 32)) * kb_stride) = pk;
        if (want_rm) *reinterpret_cast<uint32_t *>(dst_rm + j * NTH * 4) = pk;
        const int sum = __reduce_add_sync(0xffffffffu, sum4(pk));
        const int asum = __reduce_add_sync(0xffffffffu, abssum4(pk));
        if (lane == 0) {
            const int g = g4 ? (kb >> 2) : ((kb * 205) >> 10);   // kb / 5, exact for kb < 1029
            atomicAdd(&s_gsum[g], sum);
            atomicAdd(&s_gabs[g], asum);
        }
    }
    __syncthreads();
    if (warp == 0) {
        // exclusive prefix sums of the group sums (warp scan, 32 groups per pass)
        int32_t *pre = a.gpre + (int64_t) t * (a.G + 1);
        int carry = 0;
        bool safe = true;
        for (int base = 0; base < a.G; base += 32) {
            const int g = base + lane;
            const int gv = (g < a.G) ? s_gsum[g] : 0;
            safe = safe && !((g < a.G) && s_gabs[g] > 32767);
            int inc = gv;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += y;
            }
            if (g < a.G) pre[g] = carry + inc - gv;
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
        safe = __all_sync(0xffffffffu, safe);
        if (lane == 0) {
            pre[a.G] = carry;
            a.nowrap[t] = safe ? 1 : 0;
            // a token tile takes the single-drain path unless one of its tokens stamps it with this call's epoch
            if (!safe) { a.tile_unsafe[t / a.NT] = a.epoch; a.tile_unsafe[a.Tpad / 16] = a.epoch; }   // [Tpad / 16]: "some tile is unsafe"
        }
    }
}

// Rewrite q_sw from q_rm with swizzled tiles of `a.sw_tile` tokens (used when a GEMM kernel that wants the other tile height
// is forced on a workspace quantised for the default one).  One thread per 16-byte chunk.
__global__ void relayout_kernel(ActView a) {
    const int64_t chunks_per_row = a.Kpad / 16;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t) a.Tpad * chunks_per_row) return;
    const int64_t t = i / chunks_per_row;
    const int64_t k = (i % chunks_per_row) * 16;
    const uint4 v = *reinterpret_cast<const uint4 *>(a.q_rm + t * a.Kpad + k);
    const int64_t kb = k / BLOCK_K;
    const int c = (int) (k % BLOCK_K) / 16;
    int8_t *row = a.q_sw + (((t / a.sw_tile) * (a.Kpad / BLOCK_K) + kb) * a.sw_tile + (t % a.sw_tile)) * 128;
    if (!a.perm) {
        *reinterpret_cast<uint4 *>(row + ((c ^ (int) (t & 7)) << 4)) = v;
    } else {
        const uint32_t pk[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int pos = i * 32 + c * 4;
            *reinterpret_cast<uint32_t *>(row + ((((pos >> 4) ^ (int) (t & 7)) << 4) | (pos & 15))) = pk[i];
        }
    }
}

cudaError_t launch_relayout(const ActView &a, cudaStream_t s) {
    const int64_t n = (int64_t) a.Tpad * (a.Kpad / 16);
    relayout_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(a);
    return cudaGetLastError();
}

// Materialise q_rm from q_sw (inverse of the tiling / swizzle / regrouping above): only the scalar anchor kernel, the LUT
// candidate, the re-tiling fallback and the test read-back need the row-major copy, so the hot path does not write it.
__global__ void unswizzle_kernel(ActView a) {
    const int64_t chunks_per_row = a.Kpad / 16;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t) a.Tpad * chunks_per_row) return;
    const int64_t t = i / chunks_per_row;
    const int64_t k = (i % chunks_per_row) * 16;
    const int64_t kb = k / BLOCK_K;
    const int c = (int) (k % BLOCK_K) / 16;
    const int8_t *row = a.q_sw + (((t / a.sw_tile) * (a.Kpad / BLOCK_K) + kb) * a.sw_tile + (t % a.sw_tile)) * 128;
    uint4 v;
    if (!a.perm) {
        v = *reinterpret_cast<const uint4 *>(row + ((c ^ (int) (t & 7)) << 4));
    } else {
        uint32_t pk[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int pos = j * 32 + c * 4;
            pk[j] = *reinterpret_cast<const uint32_t *>(row + ((((pos >> 4) ^ (int) (t & 7)) << 4) | (pos & 15)));
        }
        v = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    }
    *reinterpret_cast<uint4 *>(a.q_rm + t * a.Kpad + k) = v;
}

cudaError_t launch_unswizzle(const ActView &a, cudaStream_t s) {
    const int64_t n = (int64_t) a.Tpad * (a.Kpad / 16);
    unswizzle_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s>>>(a);
    return cudaGetLastError();
}

template <int NTH, int CH>
static cudaError_t launch_reg(const float *x, int64_t ldx, int64_t K, int64_t T, const ActView &a, cudaStream_t s, bool pdl,
                              int early, int want_rm) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) a.Tpad);
    cfg.blockDim = dim3(NTH);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, quantize_reg_kernel<NTH, CH>, x, ldx, K, T, a, early, want_rm);
}

template <int NTH>
static cudaError_t launch_reg_n(int words, const float *x, int64_t ldx, int64_t K, int64_t T, const ActView &a, cudaStream_t s,
                                bool pdl, int early, int want_rm) {
    const int ch = (words + NTH - 1) / NTH;   // 16-byte words per thread
    // exact word counts up to 8 per thread: every word held costs 4 registers of a row that lives in registers between the
    // two passes, and the registers decide how many rows an SM has in flight (K = 8640 on 512 threads: 5 words, not 8)
    if (ch <= 1) return launch_reg<NTH, 1>(x, ldx, K, T, a, s, pdl, early, want_rm);
    if (ch <= 2) return launch_reg<NTH, 2>(x, ldx, K, T, a, s, pdl, early, want_rm);
    if (ch <= 3) return launch_reg<NTH, 3>(x, ldx, K, T, a, s, pdl, early, want_rm);
    if (ch <= 4) return launch_reg<NTH, 4>(x, ldx, K, T, a, s, pdl, early, want_rm);
    if (ch <= 5) return launch_reg<NTH, 5>(x, ldx, K, T, a, s, pdl, early, want_rm);
    if (ch <= 6) return launch_reg<NTH, 6>(x, ldx, K, T, a, s, pdl, early, want_rm);
    if (ch <= 7) return launch_reg<NTH, 7>(x, ldx, K, T, a, s, pdl, early, want_rm);
    if (ch <= 8) return launch_reg<NTH, 8>(x, ldx, K, T, a, s, pdl, early, want_rm);
    if (ch <= 16) return launch_reg<NTH, 16>(x, ldx, K, T, a, s, pdl, early, want_rm);
    return cudaErrorInvalidValue;
}

cudaError_t launch_quantize(const float *x, int64_t ldx, int64_t K, int64_t T, const ActView &a, cudaStream_t s, bool pdl,
                            bool early, bool want_rm) {
    if (a.G > 64) return cudaErrorInvalidValue;
    const bool aligned = ((reinterpret_cast<uintptr_t>(x) & 15) == 0) && (ldx % 4 == 0);
    const Knobs &kn = knobs();
    if (kn.quant_kernel != 2 && aligned && a.Kpad <= 16384) {
        const int words = (int) (a.Kpad / 4);
        // threads per token: 256; 512 for long rows of few tokens (the per-thread chain is what a small launch waits for),
        // 128 for short rows (at least a few words per thread)
        int nth = kn.quant_threads;
        // (<= 8 words per thread: 16 need 106 registers and leave 16 warps per SM, measured 18.3 vs 14.9 us at K = 14336, T = 512)
        if (nth == 0) nth = (words <= 512) ? 128 : ((words > 2048 || (a.Tpad <= 256 && words >= 2048)) ? 512 : 256);
        if (nth == 512) return launch_reg_n<512>(words, x, ldx, K, T, a, s, pdl, early, want_rm);
        if (nth == 128 && words <= 2048) return launch_reg_n<128>(words, x, ldx, K, T, a, s, pdl, early, want_rm);
        return launch_reg_n<256>(words, x, ldx, K, T, a, s, pdl, early, want_rm);
    }
    quantize_kernel<<<(unsigned) a.Tpad, QT, 0, s>>>(x, ldx, K, T, a, want_rm);
    return cudaGetLastError();
}

}  // namespace vlut
