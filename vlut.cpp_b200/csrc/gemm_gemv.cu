// gemm_gemv.cu -- VLUT_ALGO_GEMV: the whole MUL_MAT for a handful of tokens (T <= 8) on the CUDA cores, ONE launch.
//
// At T <= 4 the op is 58.7 M multiply-adds per token against 14.7 MB of packed weights: the tensor pipe is irrelevant and
// what the tcgen05 small-T kernel (gemm_tcs.cu) pays for -- TMEM allocation, decode -> MMA hand-offs, a four-accumulator
// drain, one CTA per SM, and a separate quantise kernel in front -- is all fixed latency.  This kernel is the split-K GEMV
// the memory-bound regime asks for:
//
//   * thread = one weight row of a 128-row strip; it streams its own 16 bytes per 64-k chunk of the T2 layout with
//     LDG.128 (512 contiguous bytes per warp instruction), a whole K-group (8 or 10 chunks) in flight per thread, issued
//     before griddepcontrol.wait (weights are static);
//   * mask-only unpack (as in gemm_tcs.cu): word & (0x03030303 << 2i) is four u8 codes scaled by 4^i; dp4a.u32.s32 against
//     the int8 activations accumulates bit-pair position i in its own register; the shift happens once per K-group;
//   * K is split over the S <= 8 CTAs of a thread-block cluster (whole K-groups each, so the per-group int16 wrap of the
//     reference -- ggml-cpu-quants-vlut.c:807-826 -- stays exact); the int32 partials meet in distributed shared memory;
//   * the activation quantisation (quantize_row_i8_v_tile, ggml-cpu-quants-vlut.c:952-973) is FUSED: every CTA reads the
//     f32 slice of x that belongs to its K-groups, the cluster agrees on the per-token signed maximum through DSMEM, and
//     each CTA quantises its slice into shared memory.  No workspace, no second kernel, x is read once per cluster.
//
// Small CTAs (128 threads, < 20 KB of shared memory, no TMEM): several co-resident per SM, naturally staggered.
#include "ptx.cuh"
#include "quant_math.cuh"
#include "vlut_common.cuh"

namespace vlut {
namespace gemv {

constexpr int THREADS = 128;
constexpr int MAX_GPC = 8;   // K-groups per CTA the shared-memory slice is sized for (G <= 64, S = 8)

struct Params {
    GemmArgs g;
    const float *x;
    int64_t ldx;
    int S;        // cluster size: CTAs per strip
    int G, KC;    // K-groups, 64-k chunks of the padded K
    int pdl;
};

__device__ __forceinline__ int dp4a_u8s8(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// TP: tokens per launch (T <= TP); CPG: 64-k chunks per K-group (8: I2 types, 512 k; 10: I1 types, 640 k)
template <int TP, int CPG>
__global__ void __launch_bounds__(THREADS) kernel(const __grid_constant__ Params p) {
    constexpr int GK = CPG * 64;                        // k per group
    __shared__ __align__(16) int8_t s_q[TP][MAX_GPC * GK];   // this CTA's slice of the quantised activations
    __shared__ float s_wmax[TP][THREADS / 32];
    __shared__ float s_cmax[TP];                        // this CTA's per-token maximum (read by the cluster)
    __shared__ int s_gsum[TP][MAX_GPC];
    __shared__ int32_t s_red[TP][THREADS];              // this CTA's int32 partial sums (read by the cluster)
    __shared__ float s_scale[TP];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = p.S;
    const int rank = (S > 1) ? (int) ptx::cluster_ctarank() : 0;
    const int rt_all = blockIdx.x / S;
    int ti = 0;
#pragma unroll
    for (int i = 1; i < MAX_BATCH; i++) ti += (i < p.g.n_tens && rt_all >= p.g.tens[i].strip0) ? 1 : 0;
    const TensorRef &ten = p.g.tens[ti];
    const int rt = rt_all - ten.strip0;
    const int g0 = (rank * p.G) / S, g1 = ((rank + 1) * p.G) / S;   // this CTA's K-groups
    const int ng = g1 - g0;
    const int T = (int) p.g.T;

    if (p.pdl) ptx::pdl_launch_dependents();

    // ---- weights of the first group: in flight before anything else (static data) ----
    const uint8_t *wrow = ten.w + ((int64_t) rt * p.KC) * TILE_BYTES + tid * 16;
    uint4 wv[CPG];
    auto fetch = [&](int g) {
#pragma unroll
        for (int c = 0; c < CPG; c++) {
            const int kc = g * CPG + c;
            // chunks past the padded K: weight 0 everywhere (code 1)
            wv[c] = (g < g1 && kc < p.KC) ? __ldcs(reinterpret_cast<const uint4 *>(wrow + (int64_t) kc * TILE_BYTES))
                                          : make_uint4(0x55555555u, 0x55555555u, 0x55555555u, 0x55555555u);
        }
    };
    fetch(g0);

    if (tid < TP * MAX_GPC) (&s_gsum[0][0])[tid] = 0;

    // ---- fused quantisation, pass 1: per-token signed maximum of this CTA's slice of x ----
    if (p.pdl) ptx::pdl_wait();   // x is the predecessor's output
    const int64_t k_lo = (int64_t) g0 * GK;
    const int64_t k_hi = min((int64_t) g1 * GK, p.g.K);           // real k only
    const int nvec = (int) ((k_hi - k_lo + 3) / 4);               // float4 groups (K % 4 == 0)
#pragma unroll
    for (int t = 0; t < TP; t++) {
        float mx = -INFINITY;
        if (t < T) {
            const float4 *xr = reinterpret_cast<const float4 *>(p.x + (int64_t) t * p.ldx + k_lo);
            for (int i = tid; i < nvec; i += THREADS) {
                const float4 v = __ldg(xr + i);
                mx = fmaxf(mx, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (lane == 0) s_wmax[t][warp] = mx;
    }
    __syncthreads();
    if (tid < TP) {
        float m = s_wmax[tid][0];
#pragma unroll
        for (int i = 1; i < THREADS / 32; i++) m = fmaxf(m, s_wmax[tid][i]);
        s_cmax[tid] = m;
    }
    if (S > 1) ptx::cluster_sync(); else __syncthreads();
    // every CTA of the cluster derives the same scale from the same S maxima
    if (tid < TP) {
        float m = -INFINITY;
        if (S > 1) {
            for (int r = 0; r < S; r++) {
                const uint32_t a = ptx::mapa(ptx::smem_u32(&s_cmax[tid]), (uint32_t) r);
                float v;
                asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
                m = fmaxf(m, v);
            }
        } else {
            m = s_cmax[tid];
        }
        s_wmax[tid][0] = m;
    }
    __syncthreads();

    // ---- pass 2: quantise the slice into shared memory (x re-read through L1/L2), group sums for the code offset ----
    const int slice = ng * GK;   // bytes of s_q per token (padded k hold 0)
#pragma unroll
    for (int t = 0; t < TP; t++) {
        double is = 0.0;
        float scale = 0.0f;
        if (t < T) quant_scale(s_wmax[t][0], is, scale);
        const float isf = (float) is;
        if (tid == 0) s_scale[t] = scale;
        const float4 *xr = reinterpret_cast<const float4 *>(p.x + (int64_t) t * p.ldx + k_lo);
        for (int i = tid; i < slice / 4; i += THREADS) {
            uint32_t pk = 0;
            if (t < T && i < nvec) {
                const float4 v = __ldg(xr + i);
                const int q0 = quant1(v.x, is, isf), q1 = quant1(v.y, is, isf), q2 = quant1(v.z, is, isf),
                          q3 = quant1(v.w, is, isf);
                pk = (uint32_t) (q0 & 0xff) | ((uint32_t) (q1 & 0xff) << 8) | ((uint32_t) (q2 & 0xff) << 16) |
                     ((uint32_t) (q3 & 0xff) << 24);
            }
            *reinterpret_cast<uint32_t *>(&s_q[t][4 * i]) = pk;
            // a warp covers 128 consecutive k: inside one K-group (512 and 640 are multiples of 128)
            const int sum = __reduce_add_sync(0xffffffffu, __dp4a((int) pk, 0x01010101, 0));
            if (lane == 0 && sum) atomicAdd(&s_gsum[t][(4 * i) / GK], sum);
        }
    }
    __syncthreads();

    // ---- main loop: one K-group at a time, the next group's weights in flight while this one is multiplied ----
    int32_t total[TP];
#pragma unroll
    for (int t = 0; t < TP; t++) total[t] = 0;
#pragma unroll 1
    for (int g = g0; g < g1; g++) {
        uint4 cur[CPG];
#pragma unroll
        for (int c = 0; c < CPG; c++) cur[c] = wv[c];
        fetch(g + 1);
        int acc[TP][4];
#pragma unroll
        for (int t = 0; t < TP; t++)
#pragma unroll
            for (int i = 0; i < 4; i++) acc[t][i] = 0;
        const int8_t *qg = &s_q[0][(g - g0) * GK];
#pragma unroll
        for (int c = 0; c < CPG; c++) {
            const uint32_t w4[4] = {cur[c].x, cur[c].y, cur[c].z, cur[c].w};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                // word j: k = 64 c + 16 j + 4 i + b at byte b, bit pair i  <->  activation bytes 64 c + 16 j + 4 i + b
                uint32_t m[4];
#pragma unroll
                for (int i = 0; i < 4; i++) m[i] = w4[j] & (0x03030303u << (2 * i));
#pragma unroll
                for (int t = 0; t < TP; t++) {
                    const uint4 q = *reinterpret_cast<const uint4 *>(qg + t * (MAX_GPC * GK) + c * 64 + j * 16);   // broadcast
                    acc[t][0] = dp4a_u8s8(m[0], q.x, acc[t][0]);
                    acc[t][1] = dp4a_u8s8(m[1], q.y, acc[t][1]);
                    acc[t][2] = dp4a_u8s8(m[2], q.z, acc[t][2]);
                    acc[t][3] = dp4a_u8s8(m[3], q.w, acc[t][3]);
                }
            }
        }
        // codes are w + 1: subtract sum(q) of the group; int16 wrap per K-group is part of the spec
#pragma unroll
        for (int t = 0; t < TP; t++) {
            const int v = acc[t][0] + (acc[t][1] >> 2) + (acc[t][2] >> 4) + (acc[t][3] >> 6) - s_gsum[t][g - g0];
            total[t] += wrap16(v);
        }
    }

    // ---- epilogue: partials of the S CTAs meet in distributed shared memory; CTA r finishes the tokens t = r (mod S) ----
    const int64_t n = (int64_t) rt * TILE_ROWS + tid;
    if (S == 1) {
        if (n < ten.N) {
#pragma unroll
            for (int t = 0; t < TP; t++) {
                if (t < T) {
                    float v = (float) total[t] * s_scale[t];
                    if (ten.w_scale != 1.0f) v *= ten.w_scale;
                    ten.out[(int64_t) t * ten.ldo + n] = v;
                }
            }
        }
        return;
    }
#pragma unroll
    for (int t = 0; t < TP; t++) s_red[t][tid] = total[t];
    ptx::cluster_sync();
    for (int t = rank; t < T; t += S) {
        int32_t sum = 0;
        for (int r = 0; r < S; r++)
            sum += ptx::ld_shared_cluster_s32(ptx::mapa(ptx::smem_u32(&s_red[t][tid]), (uint32_t) r));
        if (n < ten.N) {
            float v = (float) sum * s_scale[t];
            if (ten.w_scale != 1.0f) v *= ten.w_scale;
            ten.out[(int64_t) t * ten.ldo + n] = v;
        }
    }
    ptx::cluster_sync();   // nobody's shared memory goes away while a peer still reads it
}

template <int TP, int CPG>
static cudaError_t launch_tp(Params &p, cudaStream_t s) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) (p.g.tens[p.g.n_tens].strip0 * p.S));
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (p.pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (p.S > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = (unsigned) p.S;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    return cudaLaunchKernelEx(&cfg, kernel<TP, CPG>, p);
}

template <int CPG>
static cudaError_t launch_cpg(Params &p, cudaStream_t s) {
    const int T = (int) p.g.T;
    if (T <= 1) return launch_tp<1, CPG>(p, s);
    if (T <= 2) return launch_tp<2, CPG>(p, s);
    if (T <= 4) return launch_tp<4, CPG>(p, s);
    return launch_tp<8, CPG>(p, s);
}

}  // namespace gemv

bool gemm_gemv_supported(int type, int64_t K, int64_t N, int64_t T, const float *x, int64_t ldx) {
    if (!type_valid(type) || T < 1 || T > 8 || K < 4 || N < 1) return false;
    if (ceil_div(K, group_size(type)) > 64) return false;
    // vector loads of the f32 rows
    return ((reinterpret_cast<uintptr_t>(x) & 15) == 0) && (ldx % 4 == 0);
}

// the whole op (quantisation fused): x f32 [T][ldx] on the device; g.act is not used
cudaError_t launch_gemm_gemv(const GemmArgs &g, const float *x, int64_t ldx, cudaStream_t s, int *launches) {
    gemv::Params p{};
    p.g = g;
    p.x = x;
    p.ldx = ldx;
    const int gsize = group_size(g.type);
    p.G = (int) ceil_div(g.K, gsize);
    p.KC = (int) (round_up(g.K, BLOCK_K) / CHUNK_K);
    p.pdl = knobs().pdl;
    // CTAs per strip: as many as there are K-groups (at most 8: portable cluster size, and MAX_GPC groups per CTA)
    int S = p.G < 8 ? p.G : 8;
    const int forced = knobs().gemv_split;
    if (forced > 0 && forced <= 8 && forced <= p.G && ceil_div(p.G, forced) <= gemv::MAX_GPC) S = forced;
    p.S = S;
    if (launches) *launches += 1;
    return gsize == 512 ? gemv::launch_cpg<8>(p, s) : gemv::launch_cpg<10>(p, s);
}

}  // namespace vlut
