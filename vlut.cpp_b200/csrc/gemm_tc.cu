// gemm_tc.cu -- VLUT_ALGO_TC: the Vec-LUT GEMM on the 5th-generation tensor cores (tcgen05, kind::i8) of sm_100a.
//
// One kernel covers the whole token range: the UMMA N dimension is the token tile NT (16 ... 128), M = 128 weight
// rows.  The int8 weight operand never exists in HBM: packed 2-bit codes are streamed by TMA bulk copies into a
// shared-memory ring, unpacked in registers by the decode warps (shift + mask, 2 ALU ops per 4 weights) and handed
// to the tensor core either
//   TS mode: directly in TENSOR MEMORY (tcgen05.st, A operand = TMEM; lane = weight row, 32-bit column = 4 k) --
//            the decoded operand never touches shared memory, whose bandwidth would otherwise cap the kernel at
//            1/8 of the HBM rate in the small-token regime and at the MMA rate in the large-token regime; or
//   SS mode: as a 128B-swizzled K-major tile in shared memory (the conservative path, kept for comparison).
// The activation operand (B, s8, K-major, pre-swizzled by the quantise kernel) arrives with the same TMA stage.
//
// Roles (warp-specialised, 384 threads, one CTA per SM, persistent, stream-K over (output tile, K-group) units):
//   warp 0      TMA producer: per 128-k block one stage = 4 KB of T2 weights + NT x 128 B of activations
//   warp 1      MMA issuer: 4 x tcgen05.mma (K=32) per block into a double-buffered TMEM accumulator
//   warp 2      TMEM allocator
//   warps 4-7   decode: thread m owns weight row m (= TMEM lane m)
//   warps 8-11  epilogue: tcgen05.ld of the int32 accumulator, code-offset correction, int16 wrap per K-group
//               (the reference's int16 partial sums, ggml-cpu-quants-vlut.c:807-826, :80-87), fp32 scale, store
// Exactness: accumulators are drained at every K-group boundary and wrapped to int16 exactly like the reference --
// unless the quantise kernel proved that no wrap is possible for the token tile (sum|q| <= 32767 in every group),
// in which case the whole K extent is accumulated in TMEM and drained once: the result is identical.
// Tiles split across CTAs by stream-K are combined as int32 partials (order-independent, deterministic); the last
// arriving CTA applies the epilogue.
#include <stdlib.h>

#include "ptx.cuh"
#include "vlut_common.cuh"

namespace vlut {

constexpr int TC_THREADS = 384;
constexpr int TC_NA = 4;           // A-operand stages (TMEM: 32 columns each; SS mode: 16 KB of shared memory each)
constexpr int TC_MAX_STAGES = 12;

struct TcParams {
    GemmArgs g;
    int KB;      // 128-k blocks
    int G;       // K-groups
    int gblk;    // blocks per group
    int RT, TT;  // row tiles, token tiles
    int U;       // units = RT * TT * G
    int nstages;
    int stage_bytes;
    int a_smem_off;  // SS mode: offset of the A ring
    int allow_fast;
};

// ---- descriptors -------------------------------------------------------------------------------------------
// K-major, SWIZZLE_128B shared-memory operand: 8-row atoms of 1024 B (SBO), start address advanced by 32 B per
// K=32 step inside the 128-byte row.  Bits: [0,14) addr>>4, [16,30) LBO>>4 (unused, 1), [32,46) SBO>>4,
// [46,48) version = 1 (sm_100), [61,64) layout = 2 (SWIZZLE_128B).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t) ((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t) 1 << 16;
    d |= (uint64_t) (1024 >> 4) << 32;
    d |= (uint64_t) 1 << 46;
    d |= (uint64_t) 2 << 61;
    return d;
}
// instruction descriptor, kind::i8: D = s32, A = u8 (weight codes), B = s8 (activations), both K-major, M = 128
__host__ __device__ constexpr uint32_t make_idesc(int n) {
    return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (128 >> 4) << 24);
}

__device__ __forceinline__ void umma_i8_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
                 "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

struct UnitInfo {
    int tile, g, rt, tt, kb0, nblk;
};
__device__ __forceinline__ UnitInfo unit_info(const TcParams &p, int u) {
    UnitInfo i;
    i.tile = u / p.G;
    i.g = u % p.G;
    i.rt = i.tile / p.TT;
    i.tt = i.tile % p.TT;
    i.kb0 = i.g * p.gblk;
    i.nblk = min(p.gblk, p.KB - i.kb0);
    return i;
}

template <int NT, bool TS>
__global__ void __launch_bounds__(TC_THREADS, 1) gemm_tc_kernel(const __grid_constant__ TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_full[TC_MAX_STAGES], s_empty[TC_MAX_STAGES];
    __shared__ __align__(8) uint64_t s_afull[TC_NA], s_aempty[TC_NA];
    __shared__ __align__(8) uint64_t s_accfull[2], s_accempty[2];
    __shared__ uint32_t s_tmem_base;
    __shared__ int s_last;
    __shared__ int32_t s_gs[2][NT];
    __shared__ float s_sc[NT];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = p.nstages;
    // dynamic shared memory must be 1024-aligned for the 128B swizzle atoms
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) {
            ptx::mbar_init(ptx::smem_u32(&s_full[s]), 1);
            ptx::mbar_init(ptx::smem_u32(&s_empty[s]), 4 + 1);  // 4 decode warps + MMA commit
        }
        for (int j = 0; j < TC_NA; j++) {
            ptx::mbar_init(ptx::smem_u32(&s_afull[j]), 4);
            ptx::mbar_init(ptx::smem_u32(&s_aempty[j]), 1);
        }
        for (int b = 0; b < 2; b++) {
            ptx::mbar_init(ptx::smem_u32(&s_accfull[b]), 1);
            ptx::mbar_init(ptx::smem_u32(&s_accempty[b]), 4);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 2) {
        ptx::tmem_alloc(ptx::smem_u32(&s_tmem_base), 512);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = s_tmem_base;
    const uint32_t acc_col0 = 0;              // two accumulators of NT columns
    const uint32_t a_col0 = 2 * NT;           // TS mode: A ring, 32 columns per stage

    const int u0 = (int) (((int64_t) blockIdx.x * p.U) / gridDim.x);
    const int u1 = (int) (((int64_t) (blockIdx.x + 1) * p.U) / gridDim.x);

    // Is the tile of unit u on the exact (per-group drain) or the fast (single drain) path?  Uniform per tile.
    auto tile_fast = [&](int tt) -> bool {
        if (!p.allow_fast) return false;
        bool ok = true;
        for (int i = 0; i < NT; i++) ok = ok && (__ldg(p.g.act.nowrap + tt * NT + i) != 0);
        return ok;
    };

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const uint64_t pol_w = ptx::policy_evict_first();
            const uint64_t pol_a = ptx::policy_evict_last();
            uint32_t bc = 0;
            for (int u = u0; u < u1; ++u) {
                const UnitInfo ui = unit_info(p, u);
                for (int b = 0; b < ui.nblk; ++b, ++bc) {
                    const int s = bc % S;
                    const uint32_t ph = (bc / S) & 1;
                    ptx::mbar_wait(ptx::smem_u32(&s_empty[s]), ph ^ 1);
                    const uint32_t full = ptx::smem_u32(&s_full[s]);
                    ptx::mbar_arrive_expect_tx(full, BLOCK_BYTES + NT * 128);
                    const uint32_t dst = smem_base + s * p.stage_bytes;
                    const int kb = ui.kb0 + b;
                    // activations first (the MMA needs them at 1024-aligned stage base), weights behind
                    ptx::bulk_g2s_hint(dst, p.g.act.q_sw + ((int64_t) kb * p.g.act.Tpad + (int64_t) ui.tt * NT) * 128,
                                       NT * 128, full, pol_a);
                    ptx::bulk_g2s_hint(dst + NT * 128, p.g.w + ((int64_t) ui.rt * p.KB + kb) * BLOCK_BYTES, BLOCK_BYTES,
                                       full, pol_w);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(NT);
            uint32_t bc = 0, ac = 0;  // block counter, accumulator-buffer use counter
            int run_tile = -1;
            bool fast = false, open = false;
            for (int u = u0; u < u1; ++u) {
                const UnitInfo ui = unit_info(p, u);
                if (ui.tile != run_tile) {
                    run_tile = ui.tile;
                    fast = tile_fast(ui.tt);
                }
                const uint32_t ab = ac & 1;
                if (!open) {
                    ptx::mbar_wait(ptx::smem_u32(&s_accempty[ab]), ((ac >> 1) & 1) ^ 1);
                    ptx::tc_fence_after();
                }
                const uint32_t d_tmem = tmem_base + acc_col0 + ab * NT;
                for (int b = 0; b < ui.nblk; ++b, ++bc) {
                    const int s = bc % S;
                    const int j = bc % TC_NA;
                    ptx::mbar_wait(ptx::smem_u32(&s_full[s]), (bc / S) & 1);
                    ptx::mbar_wait(ptx::smem_u32(&s_afull[j]), (bc / TC_NA) & 1);
                    ptx::tc_fence_after();
                    const uint64_t b_desc = make_smem_desc(smem_base + s * p.stage_bytes);
#pragma unroll
                    for (int ks = 0; ks < 4; ks++) {
                        const uint32_t accumulate = (open || b > 0 || ks > 0) ? 1u : 0u;
                        if (TS) {
                            umma_i8_ts(d_tmem, tmem_base + a_col0 + j * 32 + ks * 8, b_desc + (uint64_t) (ks * 2), idesc,
                                       accumulate);
                        } else {
                            const uint64_t a_desc = make_smem_desc(smem_base + p.a_smem_off + j * (128 * 128));
                            ptx::umma_i8_ss(d_tmem, a_desc + (uint64_t) (ks * 2), b_desc + (uint64_t) (ks * 2), idesc,
                                            accumulate);
                        }
                    }
                    ptx::umma_commit(ptx::smem_u32(&s_empty[s]));    // stage's activation half is free again
                    ptx::umma_commit(ptx::smem_u32(&s_aempty[j]));   // A stage is free again
                }
                // exact path: every K-group is drained; fast path: drain when the tile run ends
                const bool run_end = (ui.g == p.G - 1) || (u == u1 - 1);
                if (!fast || run_end) {
                    ptx::umma_commit(ptx::smem_u32(&s_accfull[ab]));
                    ++ac;
                    open = false;
                } else {
                    open = true;
                }
            }
        }
    } else if (warp >= 4 && warp < 8) {
        // ===================== decode: packed 2-bit codes -> u8 A operand =====================
        const int dw = warp - 4;
        const int m = dw * 32 + lane;  // weight row inside the strip == TMEM lane
        uint32_t bc = 0;
        for (int u = u0; u < u1; ++u) {
            const UnitInfo ui = unit_info(p, u);
            for (int b = 0; b < ui.nblk; ++b, ++bc) {
                const int s = bc % S;
                const int j = bc % TC_NA;
                ptx::mbar_wait(ptx::smem_u32(&s_full[s]), (bc / S) & 1);
                const uint32_t w_s = smem_base + s * p.stage_bytes + NT * 128 + m * 16;
                const uint4 c0 = ptx::lds128(w_s);               // k 0..63 of the block
                const uint4 c1 = ptx::lds128(w_s + TILE_BYTES);  // k 64..127
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[s]));  // weight half of the stage consumed
                ptx::mbar_wait(ptx::smem_u32(&s_aempty[j]), ((bc / TC_NA) & 1) ^ 1);
                const uint32_t wv[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
                if (TS) {
                    ptx::tc_fence_after();
                    const uint32_t taddr = tmem_base + ((uint32_t) (dw * 32) << 16) + a_col0 + j * 32;
#pragma unroll
                    for (int ks = 0; ks < 4; ks++) {
                        uint32_t r[8];
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            r[i] = (wv[2 * ks] >> (2 * i)) & 0x03030303u;
                            r[4 + i] = (wv[2 * ks + 1] >> (2 * i)) & 0x03030303u;
                        }
                        tmem_st_32x32b_x8(taddr + ks * 8, r);
                    }
                    tmem_st_wait();
                    ptx::tc_fence_before();
                } else {
                    const uint32_t a_row = smem_base + p.a_smem_off + j * (128 * 128) + m * 128;
#pragma unroll
                    for (int c = 0; c < 8; c++) {
                        const uint32_t w = wv[c];
                        ptx::sts128(a_row + ((c ^ (m & 7)) << 4), w & 0x03030303u, (w >> 2) & 0x03030303u,
                                    (w >> 4) & 0x03030303u, (w >> 6) & 0x03030303u);
                    }
                    ptx::fence_proxy_async();
                }
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_afull[j]));
            }
        }
    } else if (warp >= 8) {
        // ===================== epilogue =====================
        const int ew = warp - 8;
        const int m = ew * 32 + lane;
        const int et = threadIdx.x - 256;  // 0..127
        int32_t tot[NT];
#pragma unroll
        for (int i = 0; i < NT; i++) tot[i] = 0;
        uint32_t ac = 0;
        int run_tile = -1, g_begin = -1;
        bool fast = false;
        for (int u = u0; u < u1; ++u) {
            const UnitInfo ui = unit_info(p, u);
            if (ui.tile != run_tile) {
                run_tile = ui.tile;
                fast = tile_fast(ui.tt);
                g_begin = ui.g;
            }
            const bool run_end = (ui.g == p.G - 1) || (u == u1 - 1);
            if (!fast || run_end) {
                const uint32_t ab = ac & 1;
                // correction term of this drain, staged through shared memory (broadcast reads below)
                if (et < NT) {
                    const int64_t t = (int64_t) ui.tt * NT + et;
                    int32_t v;
                    if (!fast) {
                        v = __ldg(p.g.act.gsum + t * p.G + ui.g);
                    } else {
                        v = 0;
                        for (int gg = g_begin; gg <= ui.g; gg++) v += __ldg(p.g.act.gsum + t * p.G + gg);
                    }
                    s_gs[ab][et] = v;
                }
                ptx::bar_sync(1, 128);
                ptx::mbar_wait(ptx::smem_u32(&s_accfull[ab]), (ac >> 1) & 1);
                ptx::tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t) (ew * 32) << 16) + acc_col0 + ab * NT;
#pragma unroll
                for (int c0 = 0; c0 < NT; c0 += 16) {
                    uint32_t v[16];
                    ptx::tmem_ld_32x32b_x16(taddr + c0, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; i++) {
                        const int32_t x = (int32_t) v[i] - s_gs[ab][c0 + i];
                        tot[c0 + i] += fast ? x : wrap16(x);
                    }
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_accempty[ab]));
                ++ac;
            }
            if (!run_end) continue;

            // ---- the tile run of this CTA is over: write out ----
            const bool complete = (g_begin == 0) && (ui.g == p.G - 1);
            const int64_t n = (int64_t) ui.rt * TILE_ROWS + m;
            const int64_t t0 = (int64_t) ui.tt * NT;
            if (et < NT) s_sc[et] = __ldg(p.g.act.scales + t0 + et);
            ptx::bar_sync(1, 128);
            if (complete) {
                if (n < p.g.N) {
#pragma unroll
                    for (int i = 0; i < NT; i++) {
                        if (t0 + i < p.g.T) {
                            float v = (float) tot[i] * s_sc[i];
                            if (p.g.w_scale != 1.0f) v *= p.g.w_scale;
                            p.g.out[(t0 + i) * p.g.ldo + n] = v;
                        }
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < NT; i++)
                    if (t0 + i < p.g.T && tot[i] != 0) atomicAdd(p.g.part + (t0 + i) * p.g.Npad + n, tot[i]);
                __threadfence();
                ptx::bar_sync(1, 128);
                if (et == 0) {
                    const int ndone = ui.g - g_begin + 1;
                    const int old = atomicAdd(p.g.counters + ui.tile, ndone);
                    s_last = (old + ndone == p.G);
                }
                ptx::bar_sync(1, 128);
                if (s_last) {
                    __threadfence();
#pragma unroll
                    for (int i = 0; i < NT; i++) {
                        if (t0 + i < p.g.T) {
                            int32_t *pp = p.g.part + (t0 + i) * p.g.Npad + n;
                            const int32_t a = __ldcg(pp);
                            *pp = 0;
                            float v = (float) a * s_sc[i];
                            if (p.g.w_scale != 1.0f) v *= p.g.w_scale;
                            if (n < p.g.N) p.g.out[(t0 + i) * p.g.ldo + n] = v;
                        }
                    }
                    if (et == 0) p.g.counters[ui.tile] = 0;
                }
            }
            ptx::bar_sync(1, 128);  // s_sc / s_last are reused by the next run
#pragma unroll
            for (int i = 0; i < NT; i++) tot[i] = 0;
            run_tile = -1;
        }
    }

    // ---- teardown ----
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 512);
    }
}

bool gemm_tc_supported(int type, int64_t K, int64_t N, int64_t T) {
    return type_valid(type) && T >= 1 && K >= 4 && N >= 1 && ceil_div(K, group_size(type)) <= 64;
}

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}

template <int NT, bool TS>
static cudaError_t launch_one(TcParams &p, int grid, cudaStream_t s) {
    const int budget = env_int("VLUT_TC_SMEM_KB", 200) * 1024;
    p.stage_bytes = NT * 128 + BLOCK_BYTES;  // multiple of 1024 for NT >= 8
    const int a_bytes = TS ? 0 : TC_NA * 128 * 128;
    int ns = (budget - a_bytes - 1024) / p.stage_bytes;
    if (ns > TC_MAX_STAGES) ns = TC_MAX_STAGES;
    const int forced = env_int("VLUT_TC_STAGES", 0);
    if (forced > 0 && forced < ns) ns = forced;
    if (ns < 2) return cudaErrorInvalidValue;
    p.nstages = ns;
    p.a_smem_off = ns * p.stage_bytes;
    const size_t smem = (size_t) ns * p.stage_bytes + a_bytes + 1024;
    auto kern = gemm_tc_kernel<NT, TS>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess) return e;
    kern<<<grid, TC_THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_gemm_tc(const GemmArgs &g, cudaStream_t s, int *launches) {
    TcParams p{};
    p.g = g;
    p.KB = (int) (g.act.Kpad / BLOCK_K);
    p.gblk = g.act.gsize / BLOCK_K;
    p.G = g.act.G;
    p.RT = (int) (g.Npad / TILE_ROWS);
    const int NT = g.act.NT;
    p.TT = (int) (g.act.Tpad / NT);
    p.U = p.RT * p.TT * p.G;
    p.allow_fast = env_int("VLUT_TC_FAST", 1);
    int grid = env_int("VLUT_TC_GRID", g.sm_count);  // override only for tests
    if (grid > p.U) grid = p.U;
    if (grid < 1) grid = 1;
    const bool ts = env_int("VLUT_TC_TS", 1) != 0;
    if (launches) *launches += 1;
    switch (NT) {
        case 16: return ts ? launch_one<16, true>(p, grid, s) : launch_one<16, false>(p, grid, s);
        case 32: return ts ? launch_one<32, true>(p, grid, s) : launch_one<32, false>(p, grid, s);
        case 64: return ts ? launch_one<64, true>(p, grid, s) : launch_one<64, false>(p, grid, s);
        case 128: return ts ? launch_one<128, true>(p, grid, s) : launch_one<128, false>(p, grid, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace vlut
