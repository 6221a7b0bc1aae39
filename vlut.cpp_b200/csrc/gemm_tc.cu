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
//   warps 0-3   epilogue: tcgen05.ld of the int32 accumulator, code-offset correction, int16 wrap per K-group
//               (the reference's int16 partial sums, ggml-cpu-quants-vlut.c:807-826, :80-87), fp32 scale, store
//   warps 4-7   decode: thread m owns weight row m (= TMEM lane m)
//   warp 8      TMA producer: per stage 4-16 KB of T2 weights + the matching activation tiles
//   warp 9      MMA issuer: 4 x tcgen05.mma (K=32) per 128-k sub-block into a double-buffered TMEM accumulator
//   warp 10     TMEM allocator
// (the warp scheduler favours the highest warp id of a sub-partition, so the latency-critical single-lane roles get
//  the top ids; measured neutral here, kept because it cannot hurt)
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
constexpr int TC_NA = 2;           // A-operand ring depth
constexpr int TC_MAX_STAGES = 12;
constexpr int SUB_K = 128;         // k per sub-block: one 128-byte swizzle row of the activation operand
constexpr int SUB_W_BYTES = 4096;  // packed weights of one sub-block (2 T2 tiles)

struct TcParams {
    GemmArgs g;
    int KB;      // 128-k sub-blocks in K
    int G;       // K-groups
    int gsub;    // sub-blocks per group (4 for I2*, 5 for I1*)
    int RT, TT;  // row tiles, token tiles
    int U;       // units = RT * TT * G
    int nstages;
    int stage_bytes;
    int a_smem_off;  // SS mode: offset of the A ring
    int allow_fast;
    long long *dbg;  // optional [grid][4 roles][8] cycle counters (VLUT_TC_DEBUG), else null
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

// The units of this CTA, walked identically by every role.  No division in the per-block path.
struct UnitWalk {
    int u, u1, G, TT, gsub, KB;
    int tile, g, rt, tt, kb0, nsub;
    __device__ __forceinline__ void init(const TcParams &p, int u0_, int u1_) {
        u = u0_; u1 = u1_; G = p.G; TT = p.TT; gsub = p.gsub; KB = p.KB;
        tile = u / G; g = u - tile * G;
        rt = tile / TT; tt = tile - rt * TT;
        refresh();
    }
    __device__ __forceinline__ void refresh() {
        kb0 = g * gsub;
        nsub = min(gsub, KB - kb0);
    }
    __device__ __forceinline__ bool valid() const { return u < u1; }
    __device__ __forceinline__ bool run_end() const { return g == G - 1 || u == u1 - 1; }
    __device__ __forceinline__ void next() {
        ++u;
        if (++g == G) {
            g = 0; ++tile;
            if (++tt == TT) { tt = 0; ++rt; }
        }
        refresh();
    }
};

struct Ring {  // stage index + phase bit, advanced without division
    int i, n;
    uint32_t ph;
    __device__ __forceinline__ void init(int n_) { i = 0; n = n_; ph = 0; }
    __device__ __forceinline__ void advance() { if (++i == n) { i = 0; ph ^= 1; } }
};

// NT: token tile (UMMA N).  SPS: 128-k sub-blocks per shared-memory stage (TMA granularity).  SPA: sub-blocks per A
// stage = per decode -> MMA hand-off (4 = 512 k = 16 MMAs per synchronisation).  TS: A operand in TMEM (else smem).
template <int NT, int SPS, int SPA, bool TS>
__global__ void __launch_bounds__(TC_THREADS, 1) gemm_tc_kernel(const __grid_constant__ TcParams p) {
    static_assert(SPA % SPS == 0, "an A stage is a whole number of shared-memory stages");
    constexpr int A_COLS = SPA * 32;                  // TMEM columns of one A stage
    constexpr int ACT_BYTES = NT * 128;               // activation operand of one sub-block
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
    if (warp == 10) {
        ptx::tmem_alloc(ptx::smem_u32(&s_tmem_base), 512);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = s_tmem_base;
    const uint32_t acc_col0 = 0;      // two accumulators of NT columns
    const uint32_t a_col0 = 2 * NT;   // TS mode: A ring, A_COLS columns per stage

    const int u0 = (int) (((int64_t) blockIdx.x * p.U) / gridDim.x);
    const int u1 = (int) (((int64_t) (blockIdx.x + 1) * p.U) / gridDim.x);

    // exact (per-group drain) or fast (single drain) path: uniform per token tile
    auto tile_fast = [&](int tt) -> bool { return p.allow_fast && (__ldg(p.g.act.tile_unsafe + tt) != p.g.act.epoch); };
    // optional instrumentation: cycles spent in each kind of wait, per role (lane 0 reports)
    long long dbgc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const bool dbg_on = p.dbg != nullptr;
    const long long t_start = dbg_on ? clock64() : 0;
    int n_ev = 0;
    int ev_role = 0;
    // timeline (CTA 0 only): time at which each wait of a role returned, 24 events per role
    auto mark = [&](int kind) {
        if (dbg_on && blockIdx.x == 0 && lane == 0 && n_ev < 24) {
            p.dbg[8192 + (ev_role * 24 + n_ev) * 2] = clock64() - t_start;
            p.dbg[8192 + (ev_role * 24 + n_ev) * 2 + 1] = kind;
            ++n_ev;
        }
    };
    auto twait = [&](uint32_t bar, uint32_t parity, int slot) {
        if (dbg_on) {
            const long long t0 = clock64();
            ptx::mbar_wait(bar, parity);
            dbgc[slot] += clock64() - t0;
            mark(slot);
        } else {
            ptx::mbar_wait(bar, parity);
        }
    };
    auto dbg_report = [&](int role) {
        if (dbg_on && lane == 0) {
            dbgc[7] = clock64() - t_start;
            for (int i = 0; i < 8; i++) p.dbg[((long long) blockIdx.x * 4 + role) * 8 + i] = dbgc[i];
        }
    };

    if (warp == 8) {
        ev_role = 0;
        // ===================== TMA producer (whole warp walks, one elected lane issues) =====================
        const uint64_t pol_w = ptx::policy_evict_first();
        const uint64_t pol_a = ptx::policy_evict_last();
        UnitWalk w;
        w.init(p, u0, u1);
        Ring st;
        st.init(S);
        for (; w.valid(); w.next()) {
            const uint8_t *wsrc = p.g.w + ((int64_t) w.rt * p.KB + w.kb0) * SUB_W_BYTES;
            // q_sw is [token tile][kb][NT][128 B]: the sub-blocks of one token tile are contiguous
            const int8_t *asrc = p.g.act.q_sw + ((int64_t) w.tt * p.KB + w.kb0) * ACT_BYTES;
            for (int sb0 = 0; sb0 < w.nsub; sb0 += SPS) {
                const int ns = min(SPS, w.nsub - sb0);
                twait(ptx::smem_u32(&s_empty[st.i]), st.ph ^ 1, 0);
                if (ptx::elect_one()) {
                    const uint32_t full = ptx::smem_u32(&s_full[st.i]);
                    ptx::mbar_arrive_expect_tx(full, ns * (SUB_W_BYTES + ACT_BYTES));
                    const uint32_t dst = smem_base + st.i * p.stage_bytes;
                    ptx::bulk_g2s_hint(dst, asrc + (int64_t) sb0 * ACT_BYTES, ns * ACT_BYTES, full, pol_a);
                    ptx::bulk_g2s_hint(dst + SPS * ACT_BYTES, wsrc + (int64_t) sb0 * SUB_W_BYTES, ns * SUB_W_BYTES, full,
                                       pol_w);
                }
                __syncwarp();
                st.advance();
            }
        }
        dbg_report(0);
    } else if (warp == 9) {
        ev_role = 1;
        // ===================== MMA issuer (whole warp walks, one elected lane issues) =====================
        constexpr uint32_t idesc = make_idesc(NT);
        UnitWalk w;
        w.init(p, u0, u1);
        Ring st, ar;
        st.init(S);
        ar.init(TC_NA);
        uint32_t ac = 0;
        int run_tile = -1;
        bool fast = false, open = false;
        for (; w.valid(); w.next()) {
            if (w.tile != run_tile) {
                run_tile = w.tile;
                fast = tile_fast(w.tt);
            }
            const uint32_t ab = ac & 1;
            if (!open) {
                twait(ptx::smem_u32(&s_accempty[ab]), ((ac >> 1) & 1) ^ 1, 2);
                ptx::tc_fence_after();
            }
            const uint32_t d_tmem = tmem_base + acc_col0 + ab * NT;
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                // the decode warps waited for every shared-memory stage of this A block before signalling it
                twait(ptx::smem_u32(&s_afull[ar.i]), ar.ph, 1);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    Ring st2 = st;
                    for (int s0 = 0; s0 < na; s0 += SPS) {
                        const int ns = min(SPS, na - s0);
                        const uint32_t stage = smem_base + st2.i * p.stage_bytes;
                        for (int sb = 0; sb < ns; sb++) {
                            const uint64_t b_desc = make_smem_desc(stage + sb * ACT_BYTES);
#pragma unroll
                            for (int ks = 0; ks < 4; ks++) {
                                const uint32_t accumulate = (open || ab0 > 0 || s0 > 0 || sb > 0 || ks > 0) ? 1u : 0u;
                                if (TS) {
                                    ptx::umma_i8_ts(d_tmem, tmem_base + a_col0 + ar.i * A_COLS + (s0 + sb) * 32 + ks * 8,
                                                    b_desc + (uint64_t) (ks * 2), idesc, accumulate);
                                } else {
                                    const uint64_t a_desc =
                                        make_smem_desc(smem_base + p.a_smem_off + (ar.i * SPA + s0 + sb) * (128 * 128));
                                    ptx::umma_i8_ss(d_tmem, a_desc + (uint64_t) (ks * 2), b_desc + (uint64_t) (ks * 2), idesc,
                                                    accumulate);
                                }
                            }
                        }
                        ptx::umma_commit(ptx::smem_u32(&s_empty[st2.i]));  // activation part of the stage is free again
                        st2.advance();
                    }
                    ptx::umma_commit(ptx::smem_u32(&s_aempty[ar.i]));  // A stage is free again
                }
                __syncwarp();
                mark(8);
                for (int s0 = 0; s0 < na; s0 += SPS) st.advance();
                ar.advance();
            }
            // exact path: every K-group is drained; fast path: drain when the tile run ends
            if (!fast || w.run_end()) {
                if (ptx::elect_one()) ptx::umma_commit(ptx::smem_u32(&s_accfull[ab]));
                __syncwarp();
                ++ac;
                open = false;
            } else {
                open = true;
            }
        }
        dbg_report(1);
    } else if (warp >= 4 && warp < 8) {
        // ===================== decode: packed 2-bit codes -> u8 A operand =====================
        const int dw = warp - 4;
        ev_role = 2;
        const int m = dw * 32 + lane;  // weight row inside the strip == TMEM lane
        UnitWalk w;
        w.init(p, u0, u1);
        Ring st, ar;
        st.init(S);
        ar.init(TC_NA);
        for (; w.valid(); w.next()) {
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                bool a_free = false;
                for (int s0 = 0; s0 < na; s0 += SPS) {
                    const int ns = min(SPS, na - s0);
                    twait(ptx::smem_u32(&s_full[st.i]), st.ph, 0);
                    const uint32_t w_s = smem_base + st.i * p.stage_bytes + SPS * ACT_BYTES + m * 16;
                    uint4 cw[2 * SPS];
#pragma unroll
                    for (int c = 0; c < 2 * SPS; c++)
                        if (c < 2 * ns) cw[c] = ptx::lds128(w_s + c * TILE_BYTES);
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[st.i]));  // weight part of the stage consumed
                    st.advance();
                    if (!a_free) {
                        twait(ptx::smem_u32(&s_aempty[ar.i]), ar.ph ^ 1, 1);
                        if (TS) ptx::tc_fence_after();
                        a_free = true;
                    }
                    if (TS) {
                        const uint32_t taddr = tmem_base + ((uint32_t) (dw * 32) << 16) + a_col0 + ar.i * A_COLS + s0 * 32;
#pragma unroll
                        for (int c = 0; c < 2 * SPS; c++) {
                            if (c < 2 * ns) {
                                const uint32_t wv[4] = {cw[c].x, cw[c].y, cw[c].z, cw[c].w};
                                uint32_t r[16];
#pragma unroll
                                for (int j = 0; j < 4; j++)
#pragma unroll
                                    for (int i = 0; i < 4; i++) r[4 * j + i] = (wv[j] >> (2 * i)) & 0x03030303u;
                                ptx::tmem_st_32x32b_x16(taddr + c * 16, r);
                            }
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 2 * SPS; c++) {
                            if (c < 2 * ns) {
                                // chunk c = k 64c..64c+63 of the stage: sub-block c/2, 16-byte swizzle chunks 4*(c%2)+j
                                const uint32_t a_row =
                                    smem_base + p.a_smem_off + (ar.i * SPA + s0 + (c >> 1)) * (128 * 128) + m * 128;
                                const uint32_t wv[4] = {cw[c].x, cw[c].y, cw[c].z, cw[c].w};
#pragma unroll
                                for (int j = 0; j < 4; j++) {
                                    const int ch = 4 * (c & 1) + j;
                                    ptx::sts128(a_row + ((ch ^ (m & 7)) << 4), wv[j] & 0x03030303u,
                                                (wv[j] >> 2) & 0x03030303u, (wv[j] >> 4) & 0x03030303u,
                                                (wv[j] >> 6) & 0x03030303u);
                                }
                            }
                        }
                    }
                }
                if (TS) {
                    if (dbg_on) {
                        const long long t0 = clock64();
                        ptx::tmem_st_wait();
                        dbgc[2] += clock64() - t0;
                    } else {
                        ptx::tmem_st_wait();
                    }
                    ptx::tc_fence_before();
                } else {
                    ptx::fence_proxy_async();
                }
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_afull[ar.i]));
                mark(9);
                ar.advance();
            }
        }
        if (dw == 0) dbg_report(2);
    } else if (warp < 4) {
        // ===================== epilogue =====================
        const int ew = warp;
        ev_role = 3;
        const int m = ew * 32 + lane;
        const int et = threadIdx.x;  // 0..127
        int32_t tot[NT];
#pragma unroll
        for (int i = 0; i < NT; i++) tot[i] = 0;
        uint32_t ac = 0;
        int run_tile = -1, g_begin = -1;
        bool fast = false, run_first = false;
        UnitWalk w;
        w.init(p, u0, u1);
        for (; w.valid(); w.next()) {
            if (w.tile != run_tile) {
                run_tile = w.tile;
                fast = tile_fast(w.tt);
                g_begin = w.g;
                run_first = (w.u == u0);
            }
            const bool run_end = w.run_end();
            if (!fast || run_end) {
                const uint32_t ab = ac & 1;
                // correction term of this drain, staged through shared memory (broadcast reads below)
                if (et < NT) {
                    const int32_t *pre = p.g.act.gpre + ((int64_t) w.tt * NT + et) * (p.G + 1);
                    s_gs[ab][et] = __ldg(pre + w.g + 1) - __ldg(pre + (fast ? g_begin : w.g));
                }
                ptx::bar_sync(1, 128);
                twait(ptx::smem_u32(&s_accfull[ab]), (ac >> 1) & 1, 0);
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
            const bool complete = (g_begin == 0) && (w.g == p.G - 1);
            const int64_t n = (int64_t) w.rt * TILE_ROWS + m;
            const int64_t t0 = (int64_t) w.tt * NT;
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
                // Split tile (stream-K): every CTA but the last to arrive parks its int32 partial tile in one of its two
                // private workspace slots (slot 0: the CTA's first run, slot 1: its last run) with plain coalesced
                // stores; the last arriver adds the peers' slots to its registers and applies the epilogue.  Integer
                // partials: the result does not depend on the arrival order.
                const int my_slot = (run_first ? 0 : 1);
                int32_t *slot = p.g.part + ((int64_t) blockIdx.x * 2 + my_slot) * (NT * 128);
#pragma unroll
                for (int i = 0; i < NT; i++) __stcg(slot + i * 128 + m, tot[i]);
                __threadfence();
                ptx::bar_sync(1, 128);
                if (et == 0) {
                    const int ndone = w.g - g_begin + 1;
                    const int old = atomicAdd(p.g.counters + w.tile, ndone);
                    s_last = (old + ndone == p.G);
                }
                ptx::bar_sync(1, 128);
                if (s_last) {
                    __threadfence();
                    // contributors: the CTAs whose unit range intersects this tile
                    const long long ua = (long long) w.tile * p.G, ub = ua + p.G - 1;
                    const int b_lo = (int) (((ua + 1) * gridDim.x - 1) / p.U);
                    const int b_hi = (int) (((ub + 1) * gridDim.x - 1) / p.U);
                    for (int b = b_lo; b <= b_hi; b++) {
                        if (b == (int) blockIdx.x) continue;
                        const int b_u0 = (int) (((long long) b * p.U) / gridDim.x);
                        const int b_slot = (b_u0 / p.G == w.tile) ? 0 : 1;
                        const int32_t *ps = p.g.part + ((int64_t) b * 2 + b_slot) * (NT * 128) + m;
#pragma unroll
                        for (int c0 = 0; c0 < NT; c0 += 16) {
                            int32_t a[16];
#pragma unroll
                            for (int i = 0; i < 16; i++) a[i] = __ldcg(ps + (c0 + i) * 128);
#pragma unroll
                            for (int i = 0; i < 16; i++) tot[c0 + i] += a[i];
                        }
                    }
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
                    if (et == 0) p.g.counters[w.tile] = 0;
                }
            }
            ptx::bar_sync(1, 128);  // s_sc / s_last are reused by the next run
#pragma unroll
            for (int i = 0; i < NT; i++) tot[i] = 0;
            run_tile = -1;
        }
        if (ew == 0) dbg_report(3);
    }

    // ---- teardown ----
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 10) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 512);
    }
}

bool gemm_tc_supported(int type, int64_t K, int64_t N, int64_t T) {
    return type_valid(type) && T >= 1 && K >= 4 && N >= 1 && ceil_div(K, group_size(type)) <= 64;
}

constexpr int DBG_WORDS = 256 * 4 * 8 + 4 * 24 * 2;
static long long *g_dbg_buf = nullptr;

// zeroed debug buffer shared by the kernels' optional instrumentation; word 0 = number of timeline events to skip
long long *debug_buffer(cudaStream_t s, int skip) {
    if (!g_dbg_buf) cudaMalloc(&g_dbg_buf, DBG_WORDS * sizeof(long long));
    cudaMemsetAsync(g_dbg_buf, 0, DBG_WORDS * sizeof(long long), s);
    const long long sk = skip;
    cudaMemcpyAsync(g_dbg_buf, &sk, sizeof(sk), cudaMemcpyHostToDevice, s);
    return g_dbg_buf;
}

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}

template <int NT, int SPS, int SPA, bool TS>
static cudaError_t launch_one(TcParams &p, int grid, cudaStream_t s) {
    const int budget = env_int("VLUT_TC_SMEM_KB", 200) * 1024;
    p.stage_bytes = SPS * (NT * 128 + SUB_W_BYTES);  // multiple of 1024
    const int a_bytes = TS ? 0 : TC_NA * SPA * 128 * 128;
    int ns = (budget - a_bytes - 1024) / p.stage_bytes;
    if (ns > TC_MAX_STAGES) ns = TC_MAX_STAGES;
    int forced = env_int("VLUT_TC_STAGES", 0);
    if (forced > 0 && forced < SPA / SPS) forced = SPA / SPS;
    if (forced > 0 && forced < ns) ns = forced;
    if (ns < SPA / SPS) return cudaErrorInvalidValue;  // all shared-memory stages of one A block are live together
    p.nstages = ns;
    p.a_smem_off = ns * p.stage_bytes;
    const size_t smem = (size_t) ns * p.stage_bytes + a_bytes + 1024;
    auto kern = gemm_tc_kernel<NT, SPS, SPA, TS>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess) return e;
    kern<<<grid, TC_THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_gemm_tc(const GemmArgs &g, cudaStream_t s, int *launches) {
    if (g.act.sw_tile != g.act.NT) return cudaErrorInvalidValue;  // wants full-height swizzled activation tiles
    TcParams p{};
    p.g = g;
    p.KB = (int) (g.act.Kpad / SUB_K);
    p.gsub = g.act.gsize / SUB_K;
    p.G = g.act.G;
    p.RT = (int) (g.Npad / TILE_ROWS);
    const int NT = g.act.NT;
    p.TT = (int) (g.act.Tpad / NT);
    p.U = p.RT * p.TT * p.G;
    p.allow_fast = env_int("VLUT_TC_FAST", 1);
    int grid = g.sm_count;
    // when the output tiles alone (nearly) fill the machine, give every CTA whole tiles: no split-tile fix-up
    const int tiles = p.RT * p.TT;
    if (tiles <= g.sm_count && tiles * 10 >= g.sm_count * 7) grid = tiles;
    grid = env_int("VLUT_TC_GRID", grid);  // override only for tests
    if (grid > p.U) grid = p.U;
    if (grid < 1) grid = 1;
    const bool ts = env_int("VLUT_TC_TS", 1) != 0;
    if (launches) *launches += 1;
    if (env_int("VLUT_TC_DEBUG", 0)) {
        if (!g_dbg_buf) cudaMalloc(&g_dbg_buf, DBG_WORDS * sizeof(long long));
        cudaMemsetAsync(g_dbg_buf, 0, DBG_WORDS * sizeof(long long), s);
        p.dbg = g_dbg_buf;
    }
    switch (NT) {
        case 16: return ts ? launch_one<16, 4, 4, true>(p, grid, s) : launch_one<16, 2, 2, false>(p, grid, s);
        case 32: return ts ? launch_one<32, 4, 4, true>(p, grid, s) : launch_one<32, 2, 2, false>(p, grid, s);
        case 64: return ts ? launch_one<64, 2, 4, true>(p, grid, s) : launch_one<64, 1, 2, false>(p, grid, s);
        case 128: return ts ? launch_one<128, 2, 4, true>(p, grid, s) : launch_one<128, 1, 2, false>(p, grid, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace vlut

// test/profiling hook: cycle counters of the last launch made with VLUT_TC_DEBUG=1, [256 CTAs][4 roles][8] int64
extern "C" int vlut_debug_tc_counters(long long *out_host, int n_words) {
    if (!out_host || n_words <= 0 || n_words > vlut::DBG_WORDS) return VLUT_ERR_INVALID;
    if (!vlut::g_dbg_buf) return VLUT_ERR_UNSUPPORTED;
    if (cudaDeviceSynchronize() != cudaSuccess) return VLUT_ERR_CUDA;
    if (cudaMemcpy(out_host, vlut::g_dbg_buf, (size_t) n_words * sizeof(long long), cudaMemcpyDeviceToHost) != cudaSuccess)
        return VLUT_ERR_CUDA;
    return VLUT_OK;
}
