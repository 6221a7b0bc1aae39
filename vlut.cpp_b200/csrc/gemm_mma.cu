// gemm_mma.cu -- VLUT_ALGO_MMA: the small-token-count (HBM-bound) Vec-LUT GEMM for sm_100a.
//
// Design (DESIGN.md "Kernel S"):
//   * persistent grid, one CTA per SM, stream-K: the (row-tile, K-group) units of the whole matrix are split
//     evenly over the CTAs, so all 148 SMs stream the same number of packed-weight bytes whatever N is;
//   * a single producer thread streams each unit -- 16-20 KB of contiguous T2 weight tiles plus the matching
//     activation fragments -- with TMA bulk copies (cp.async.bulk, SASS UBLKCP) into a ring of shared-memory
//     stages guarded by full/empty mbarriers: every byte the SM will consume can be in flight at once;
//   * consumer warps read their 16-byte lane entries conflict-free (LDS.128), unpack the 2-bit codes in
//     registers (shift+mask: 7 ALU ops per 16 weights) and feed mma.sync.m16n8k32 (s8 activations x u8 codes,
//     s32 accumulators) with the token rows as M and 8 weight rows as N -- no decoded weight ever touches
//     shared memory, which would cost 8x the HBM rate in shared-memory bandwidth;
//   * at every K-group boundary the int32 partial is corrected for the +1 code offset (gsum), wrapped to int16
//     like the reference's int16 accumulators (ggml-cpu-quants-vlut.c:807-826, :80-87) and added to the total;
//   * row tiles whose groups were all done by this CTA are scaled and stored directly; tiles split across
//     CTAs are combined through int32 atomics in a self-cleaning workspace, the last arriving CTA applies
//     the fp32 epilogue (ggml-cpu-quants-vlut.c:1252-1261).  Integer partials make the result independent of
//     the arrival order: bit-exact and deterministic.
#include <stdlib.h>

#include "ptx.cuh"
#include "vlut_common.cuh"

namespace vlut {

struct MmaParams {
    GemmArgs g;
    int KB;       // 128-wide k blocks (Kpad / 128)
    int G;        // K-groups
    int gblk;     // k blocks per group (4 for I2*, 5 for I1*)
    int RT;       // 128-row tiles
    int U;        // units = RT * G
    int MTall;    // m16 token tiles present in q_frag
    int nstages;
    int stage_bytes;
    int stage_w_bytes;
    int even_ldo;
};

template <int MT, int CH>
__global__ void __launch_bounds__((8 / CH + 1) * 32, 1) gemm_mma_kernel(const __grid_constant__ MmaParams p) {
    constexpr int NW = 8 / CH;  // consumer warps
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t s_full[8];
    __shared__ __align__(8) uint64_t s_empty[8];
    __shared__ int s_last;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = p.nstages;
    const uint32_t smem_base = ptx::smem_u32(smem);

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) {
            ptx::mbar_init(ptx::smem_u32(&s_full[s]), 1);
            ptx::mbar_init(ptx::smem_u32(&s_empty[s]), NW);
        }
        ptx::fence_barrier_init();
    }
    __syncthreads();

    // stream-K partition of the units
    const int u0 = (int) (((int64_t) blockIdx.x * p.U) / gridDim.x);
    const int u1 = (int) (((int64_t) (blockIdx.x + 1) * p.U) / gridDim.x);
    const int MTall = p.MTall;

    if (warp == 0) {
        // ===== producer: one thread issues all TMA bulk copies =====
        if (lane == 0) {
            const uint64_t pol_w = ptx::policy_evict_first();
            const uint64_t pol_a = ptx::policy_evict_last();
            uint32_t cnt = 0;
            for (int u = u0; u < u1; ++u, ++cnt) {
                const int s = cnt % S;
                const uint32_t ph = (cnt / S) & 1;
                ptx::mbar_wait(ptx::smem_u32(&s_empty[s]), ph ^ 1);
                const int rt = u / p.G, g = u % p.G;
                const int kb0 = g * p.gblk;
                const int nblk = min(p.gblk, p.KB - kb0);
                const uint32_t wbytes = (uint32_t) nblk * TILE_BYTES;
                const uint32_t abytes = (uint32_t) nblk * MTall * 2048;
                const uint32_t full = ptx::smem_u32(&s_full[s]);
                ptx::mbar_arrive_expect_tx(full, wbytes + abytes);
                const uint32_t dst = smem_base + s * p.stage_bytes;
                ptx::bulk_g2s_hint(dst, p.g.w + ((int64_t) rt * p.KB + kb0) * TILE_BYTES, wbytes, full, pol_w);
                ptx::bulk_g2s_hint(dst + p.stage_w_bytes, p.g.act.q_frag + (int64_t) kb0 * MTall * 2048, abytes, full,
                                   pol_a);
            }
        }
        return;
    }

    // ===== consumers =====
    const int cw = warp - 1;
    const int gq = lane >> 2, tig = lane & 3;
    int32_t tot[MT][2 * CH][4];
#pragma unroll
    for (int a = 0; a < MT; a++)
#pragma unroll
        for (int b = 0; b < 2 * CH; b++)
#pragma unroll
            for (int c = 0; c < 4; c++) tot[a][b][c] = 0;

    int g_begin = -1;
    uint32_t cnt = 0;
    for (int u = u0; u < u1; ++u, ++cnt) {
        const int s = cnt % S;
        const uint32_t ph = (cnt / S) & 1;
        const int rt = u / p.G, g = u % p.G;
        const int kb0 = g * p.gblk;
        const int nblk = min(p.gblk, p.KB - kb0);
        if (g_begin < 0) g_begin = g;

        // code-offset correction of this group, fetched while the stage is still in flight
        int32_t gs[MT][2];
#pragma unroll
        for (int mt = 0; mt < MT; mt++)
#pragma unroll
            for (int rh = 0; rh < 2; rh++)
                gs[mt][rh] = (mt < MTall) ? __ldg(p.g.act.gsum + (int64_t) (16 * mt + gq + 8 * rh) * p.G + g) : 0;

        int32_t acc[MT][2 * CH][4];
#pragma unroll
        for (int a = 0; a < MT; a++)
#pragma unroll
            for (int b = 0; b < 2 * CH; b++)
#pragma unroll
                for (int c = 0; c < 4; c++) acc[a][b][c] = 0;

        ptx::mbar_wait(ptx::smem_u32(&s_full[s]), ph);
        const uint32_t w_s = smem_base + s * p.stage_bytes + (cw * CH) * CHUNK_BYTES + lane * 16;
        const uint32_t a_s = smem_base + s * p.stage_bytes + p.stage_w_bytes + lane * 16;

        for (int b = 0; b < nblk; b++) {
            uint4 wq[CH];
#pragma unroll
            for (int c = 0; c < CH; c++) wq[c] = ptx::lds128(w_s + b * TILE_BYTES + c * CHUNK_BYTES);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                uint4 af[MT];
#pragma unroll
                for (int mt = 0; mt < MT; mt++)
                    if (mt < MTall) af[mt] = ptx::lds128(a_s + ((b * MTall + mt) * 4 + i) * 512);
#pragma unroll
                for (int c = 0; c < CH; c++) {
                    const uint32_t b00 = (wq[c].x >> (2 * i)) & 0x03030303u;  // rg 0, h 0
                    const uint32_t b01 = (wq[c].y >> (2 * i)) & 0x03030303u;  // rg 0, h 1
                    const uint32_t b10 = (wq[c].z >> (2 * i)) & 0x03030303u;  // rg 1, h 0
                    const uint32_t b11 = (wq[c].w >> (2 * i)) & 0x03030303u;  // rg 1, h 1
#pragma unroll
                    for (int mt = 0; mt < MT; mt++)
                        if (mt < MTall) {
                            ptx::mma_s8u8(acc[mt][2 * c + 0], af[mt], b00, b01);
                            ptx::mma_s8u8(acc[mt][2 * c + 1], af[mt], b10, b11);
                        }
                }
            }
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[s]));

        // K-group boundary: undo the +1 code offset, wrap to int16, accumulate
#pragma unroll
        for (int mt = 0; mt < MT; mt++)
#pragma unroll
            for (int jj = 0; jj < 2 * CH; jj++)
#pragma unroll
                for (int r = 0; r < 4; r++) tot[mt][jj][r] += wrap16(acc[mt][jj][r] - gs[mt][r >> 1]);

        const bool strip_end = (g == p.G - 1) || (u == u1 - 1);
        if (!strip_end) continue;

        const bool complete = (g_begin == 0) && (g == p.G - 1);
        const int64_t nbase = (int64_t) rt * TILE_ROWS + (cw * CH) * CHUNK_ROWS + 2 * tig;
        if (complete) {
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
                if (mt >= MTall) continue;
#pragma unroll
                for (int rh = 0; rh < 2; rh++) {
                    const int64_t t = 16 * mt + gq + 8 * rh;
                    if (t >= p.g.T) continue;
                    const float sc = __ldg(p.g.act.scales + t);
#pragma unroll
                    for (int jj = 0; jj < 2 * CH; jj++) {
                        const int64_t n = nbase + (jj >> 1) * CHUNK_ROWS + (jj & 1) * 8;
                        float v0 = (float) tot[mt][jj][2 * rh] * sc;
                        float v1 = (float) tot[mt][jj][2 * rh + 1] * sc;
                        if (p.g.w_scale != 1.0f) { v0 *= p.g.w_scale; v1 *= p.g.w_scale; }
                        float *o = p.g.out + t * p.g.ldo + n;
                        if (p.even_ldo && n + 1 < p.g.N) {
                            *reinterpret_cast<float2 *>(o) = make_float2(v0, v1);
                        } else {
                            if (n < p.g.N) o[0] = v0;
                            if (n + 1 < p.g.N) o[1] = v1;
                        }
                    }
                }
            }
        } else {
            // split tile: integer partials through the workspace, last arriver finishes
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
                if (mt >= MTall) continue;
#pragma unroll
                for (int rh = 0; rh < 2; rh++) {
                    const int64_t t = 16 * mt + gq + 8 * rh;
                    if (t >= p.g.T) continue;
#pragma unroll
                    for (int jj = 0; jj < 2 * CH; jj++) {
                        const int64_t n = nbase + (jj >> 1) * CHUNK_ROWS + (jj & 1) * 8;
                        int32_t *pp = p.g.part + t * p.g.Npad + n;
                        if (tot[mt][jj][2 * rh] != 0) atomicAdd(pp, tot[mt][jj][2 * rh]);
                        if (tot[mt][jj][2 * rh + 1] != 0) atomicAdd(pp + 1, tot[mt][jj][2 * rh + 1]);
                    }
                }
            }
            __threadfence();
            ptx::bar_sync(1, NW * 32);
            if (cw == 0 && lane == 0) {
                const int ndone = g - g_begin + 1;
                const int old = atomicAdd(p.g.counters + rt, ndone);
                s_last = (old + ndone == p.G);
            }
            ptx::bar_sync(1, NW * 32);
            if (s_last) {
                __threadfence();
#pragma unroll
                for (int mt = 0; mt < MT; mt++) {
                    if (mt >= MTall) continue;
#pragma unroll
                    for (int rh = 0; rh < 2; rh++) {
                        const int64_t t = 16 * mt + gq + 8 * rh;
                        if (t >= p.g.T) continue;
                        const float sc = __ldg(p.g.act.scales + t);
#pragma unroll
                        for (int jj = 0; jj < 2 * CH; jj++) {
                            const int64_t n = nbase + (jj >> 1) * CHUNK_ROWS + (jj & 1) * 8;
                            int32_t *pp = p.g.part + t * p.g.Npad + n;
                            const int2 v = __ldcg(reinterpret_cast<const int2 *>(pp));
                            *reinterpret_cast<int2 *>(pp) = make_int2(0, 0);
                            float v0 = (float) v.x * sc, v1 = (float) v.y * sc;
                            if (p.g.w_scale != 1.0f) { v0 *= p.g.w_scale; v1 *= p.g.w_scale; }
                            float *o = p.g.out + t * p.g.ldo + n;
                            if (n < p.g.N) o[0] = v0;
                            if (n + 1 < p.g.N) o[1] = v1;
                        }
                    }
                }
                if (cw == 0 && lane == 0) p.g.counters[rt] = 0;
            }
            ptx::bar_sync(1, NW * 32);  // s_last is reused by the next split tile
        }
#pragma unroll
        for (int a = 0; a < MT; a++)
#pragma unroll
            for (int b = 0; b < 2 * CH; b++)
#pragma unroll
                for (int c = 0; c < 4; c++) tot[a][b][c] = 0;
        g_begin = -1;
    }
}

bool gemm_mma_supported(int type, int64_t K, int64_t N, int64_t T) {
    return type_valid(type) && T >= 1 && T <= 64 && K >= 4 && N >= 1;
}

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}

template <int MT, int CH>
static cudaError_t launch_one(const MmaParams &p, int grid, size_t smem, cudaStream_t s) {
    auto kern = gemm_mma_kernel<MT, CH>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess) return e;
    kern<<<grid, (8 / CH + 1) * 32, smem, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_gemm_mma(const GemmArgs &g, cudaStream_t s, int *launches) {
    MmaParams p;
    p.g = g;
    p.KB = (int) (g.act.Kpad / TILE_K);
    p.gblk = g.act.gsize / TILE_K;
    p.G = g.act.G;
    p.RT = (int) (g.Npad / TILE_ROWS);
    p.U = p.RT * p.G;
    p.MTall = (int) (g.act.Tpad / 16);
    p.even_ldo = (g.ldo % 2 == 0) && ((reinterpret_cast<uintptr_t>(g.out) & 7) == 0);
    p.stage_w_bytes = p.gblk * TILE_BYTES;
    p.stage_bytes = p.stage_w_bytes + p.gblk * p.MTall * 2048;
    const int budget = env_int("VLUT_MMA_SMEM_KB", 200) * 1024;
    p.nstages = budget / p.stage_bytes;
    if (p.nstages > 8) p.nstages = 8;
    if (p.nstages < 2) return cudaErrorInvalidValue;
    const size_t smem = (size_t) p.nstages * p.stage_bytes;
    int grid = g.sm_count * env_int("VLUT_MMA_CTAS_PER_SM", 1);
    if (grid > p.U) grid = p.U;
    const int ch = env_int("VLUT_MMA_CH", 2);
    if (launches) *launches += 1;
    if (p.MTall == 1) return ch == 1 ? launch_one<1, 1>(p, grid, smem, s) : launch_one<1, 2>(p, grid, smem, s);
    if (p.MTall == 2) return ch == 1 ? launch_one<2, 1>(p, grid, smem, s) : launch_one<2, 2>(p, grid, smem, s);
    if (p.MTall <= 4) return ch == 1 ? launch_one<4, 1>(p, grid, smem, s) : launch_one<4, 2>(p, grid, smem, s);
    return cudaErrorInvalidValue;
}

}  // namespace vlut
