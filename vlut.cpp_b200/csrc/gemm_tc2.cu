// gemm_tc2.cu -- VLUT_ALGO_TC: the Vec-LUT GEMM on tcgen05 (kind::i8), two interleaved decode+issue pipelines per CTA.
//
// Same algorithm, data layouts and exactness contract as gemm_tc.cu (see its header and DESIGN.md section 4.1); what
// changes is the intra-CTA organisation.  Cycle counters showed the single-pipeline kernel to be LATENCY-bound: per
// 512-k block the decode warps and the MMA issuer each spend ~1000 clk in serial instruction latency (mbarrier
// try_wait, LDS, tcgen05.st + wait::st, uniform-datapath moves) while the tensor pipe needs 288 clk (N=16) ... 1024
// clk (N=128) for the 16 MMAs of that block.  Here two independent pipelines alternate over the (tile, K-group)
// units of the CTA, so that one pipeline's latencies hide behind the other's work:
//
//   warps 0-3    epilogue (TMEM lane quadrant = warp)           warp 12   TMA producer (shared stage ring)
//   warps 4-7    decode set 0  -> A stage 0 (TMEM cols)         warp 13   MMA issuer 0 -> accumulators of pipeline 0
//   warps 8-11   decode set 1  -> A stage 1                     warp 14   MMA issuer 1 -> accumulators of pipeline 1
//                                                               warp 15   TMEM allocator
//   pipeline d owns the units u with (u - u0) % 2 == d; a unit = one K-group (512 / 640 k) of one 128 x NT tile.
//   Each pipeline has its OWN ring of shared-memory stages (the producer fills the ring of the unit's owner): a
//   consumer that skipped phases of a shared ring would alias mbarrier phase parities.
//
// The epilogue is register-light: accumulators are consumed 16 columns at a time (tcgen05.ld.x16); in the exact path
// the per-K-group int16 wrap totals live in shared memory, in the fast path (no wrap possible, proven by the quantise
// kernel) nothing is kept at all -- the two pipelines' accumulators are summed as they are read.  Every variant fits
// the 128 registers/thread that 512 threads allow.
#include <stdlib.h>

#include "ptx.cuh"
#include "vlut_common.cuh"

namespace vlut {

namespace tc2 {

constexpr int THREADS = 512;
constexpr int SPA = 4;             // 128-k sub-blocks per A stage / decode->MMA hand-off
constexpr int A_COLS = SPA * 32;   // TMEM columns of one A stage
constexpr int MAX_STAGES = 12;
constexpr int SUB_W_BYTES = 4096;

struct Params {
    GemmArgs g;
    int KB, G, gsub, RT, TT, U;
    int nstages, stage_bytes, tots_off;
    int allow_fast;
    int ablate;  // profiling only (VLUT_TC_ABLATE): 1 skip decode ALU, 2 skip tcgen05.st, 4 skip tcgen05.mma -> WRONG results
};

__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t) ((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t) 1 << 16;
    d |= (uint64_t) (1024 >> 4) << 32;
    d |= (uint64_t) 1 << 46;
    d |= (uint64_t) 2 << 61;
    return d;
}
__host__ __device__ constexpr uint32_t make_idesc(int n) {
    return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (128 >> 4) << 24);
}

struct Walk {  // the units of this CTA, walked identically by every role; no division in the per-block path
    int u, u0, u1, G, TT, gsub, KB;
    int tile, g, rt, tt, kb0, nsub;
    __device__ __forceinline__ void init(const Params &p, int u0_, int u1_) {
        u = u0 = u0_; u1 = u1_; G = p.G; TT = p.TT; gsub = p.gsub; KB = p.KB;
        tile = u / G; g = u - tile * G;
        rt = tile / TT; tt = tile - rt * TT;
        refresh();
    }
    __device__ __forceinline__ void refresh() { kb0 = g * gsub; nsub = min(gsub, KB - kb0); }
    __device__ __forceinline__ bool valid() const { return u < u1; }
    __device__ __forceinline__ int owner() const { return (u - u0) & 1; }
    __device__ __forceinline__ int run_first() const { return max(u0, u - g); }            // first unit of this tile run
    __device__ __forceinline__ int run_last() const { return min(u1 - 1, u - g + G - 1); }  // last unit of this tile run
    __device__ __forceinline__ void next() {
        ++u;
        if (++g == G) { g = 0; ++tile; if (++tt == TT) { tt = 0; ++rt; } }
        refresh();
    }
};

struct Ring {
    int i, n;
    uint32_t ph;
    __device__ __forceinline__ void init(int n_) { i = 0; n = n_; ph = 0; }
    __device__ __forceinline__ void advance() { if (++i == n) { i = 0; ph ^= 1; } }
};

template <int NT, int SPS>
__global__ void __launch_bounds__(THREADS, 1) kernel(const __grid_constant__ Params p) {
    static_assert(SPA % SPS == 0, "an A stage is a whole number of shared-memory stages");
    constexpr int NACC = (NT == 128) ? 1 : 2;  // accumulator buffers per pipeline
    constexpr int ACT_BYTES = NT * 128;
    constexpr uint32_t A_COL0 = 2 * NACC * NT;
    static_assert(A_COL0 + 2 * A_COLS <= 512, "TMEM budget");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_full[2][MAX_STAGES / 2], s_empty[2][MAX_STAGES / 2];
    __shared__ __align__(8) uint64_t s_afull[2], s_aempty[2];
    __shared__ __align__(8) uint64_t s_accfull[2][2], s_accempty[2][2];
    __shared__ uint32_t s_tmem_base;
    __shared__ int s_last;
    __shared__ int32_t s_gs[NT];
    __shared__ float s_sc[NT];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = p.nstages / 2;  // stages per pipeline
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;

    if (threadIdx.x == 0) {
        for (int d = 0; d < 2; d++) {
            for (int s = 0; s < S; s++) {
                ptx::mbar_init(ptx::smem_u32(&s_full[d][s]), 1);
                ptx::mbar_init(ptx::smem_u32(&s_empty[d][s]), 4 + 1);  // the pipeline's 4 decode warps + its MMA commit
            }
            ptx::mbar_init(ptx::smem_u32(&s_afull[d]), 4);
            ptx::mbar_init(ptx::smem_u32(&s_aempty[d]), 1);
            for (int b = 0; b < 2; b++) {
                ptx::mbar_init(ptx::smem_u32(&s_accfull[d][b]), 1);
                ptx::mbar_init(ptx::smem_u32(&s_accempty[d][b]), 4);
            }
        }
        ptx::fence_barrier_init();
    }
    if (warp == 15) {
        ptx::tmem_alloc(ptx::smem_u32(&s_tmem_base), 512);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = s_tmem_base;

    const int u0 = (int) (((int64_t) blockIdx.x * p.U) / gridDim.x);
    const int u1 = (int) (((int64_t) (blockIdx.x + 1) * p.U) / gridDim.x);
    auto tile_fast = [&](int tt) -> bool { return p.allow_fast && (__ldg(p.g.act.tile_fast + tt) != 0); };

    if (warp == 12) {
        // ===================== TMA producer =====================
        const uint64_t pol_w = ptx::policy_evict_first();
        const uint64_t pol_a = ptx::policy_evict_last();
        Walk w;
        w.init(p, u0, u1);
        Ring rings[2];
        rings[0].init(S);
        rings[1].init(S);
        for (; w.valid(); w.next()) {
            const int d = w.owner();
            Ring &st = rings[d];
            const uint8_t *wsrc = p.g.w + ((int64_t) w.rt * p.KB + w.kb0) * SUB_W_BYTES;
            const int8_t *asrc = p.g.act.q_sw + ((int64_t) w.tt * p.KB + w.kb0) * ACT_BYTES;
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                for (int s0 = 0; s0 < na; s0 += SPS) {
                    const int ns = min(SPS, na - s0);
                    ptx::mbar_wait(ptx::smem_u32(&s_empty[d][st.i]), st.ph ^ 1);
                    if (ptx::elect_one()) {
                        const uint32_t full = ptx::smem_u32(&s_full[d][st.i]);
                        ptx::mbar_arrive_expect_tx(full, ns * (SUB_W_BYTES + ACT_BYTES));
                        const uint32_t dst = smem_base + (d * S + st.i) * p.stage_bytes;
                        ptx::bulk_g2s_hint(dst, asrc + (int64_t) (ab0 + s0) * ACT_BYTES, ns * ACT_BYTES, full, pol_a);
                        ptx::bulk_g2s_hint(dst + SPS * ACT_BYTES, wsrc + (int64_t) (ab0 + s0) * SUB_W_BYTES, ns * SUB_W_BYTES,
                                           full, pol_w);
                    }
                    __syncwarp();
                    st.advance();
                }
            }
        }
    } else if (warp == 13 || warp == 14) {
        // ===================== MMA issuer of pipeline d =====================
        const int d = warp - 13;
        constexpr uint32_t idesc = make_idesc(NT);
        Walk w;
        w.init(p, u0, u1);
        Ring st;
        st.init(S);
        uint32_t a_ph = 0;    // phase of this pipeline's (single) A stage
        uint32_t ndrain = 0;  // drains committed by this pipeline so far
        int run_tile = -1;
        bool fast = false, open = false;
        for (; w.valid(); w.next()) {
            if (w.owner() != d) continue;
            if (w.tile != run_tile) {
                run_tile = w.tile;
                fast = tile_fast(w.tt);
            }
            const uint32_t ab = ndrain % NACC;
            if (!open) {
                ptx::mbar_wait(ptx::smem_u32(&s_accempty[d][ab]), ((ndrain / NACC) & 1) ^ 1);
                ptx::tc_fence_after();
            }
            const uint32_t d_tmem = tmem_base + (d * NACC + ab) * NT;
            const uint32_t a_tmem = tmem_base + A_COL0 + d * A_COLS;
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                ptx::mbar_wait(ptx::smem_u32(&s_afull[d]), a_ph);
                a_ph ^= 1;
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    Ring st2 = st;
                    for (int s0 = 0; s0 < na; s0 += SPS) {
                        const int ns = min(SPS, na - s0);
                        const uint32_t stage = smem_base + (d * S + st2.i) * p.stage_bytes;
                        for (int sb = 0; sb < ns; sb++) {
                            const uint64_t b_desc = make_smem_desc(stage + sb * ACT_BYTES);
#pragma unroll
                            for (int ks = 0; ks < 4; ks++) {
                                const uint32_t accumulate = (open || ab0 > 0 || s0 > 0 || sb > 0 || ks > 0) ? 1u : 0u;
                                if (!(p.ablate & 4)) ptx::umma_i8_ts(d_tmem, a_tmem + (s0 + sb) * 32 + ks * 8, b_desc + (uint64_t) (ks * 2), idesc,
                                                accumulate);
                            }
                        }
                        ptx::umma_commit(ptx::smem_u32(&s_empty[d][st2.i]));
                        st2.advance();
                    }
                    ptx::umma_commit(ptx::smem_u32(&s_aempty[d]));
                }
                __syncwarp();
                for (int s0 = 0; s0 < na; s0 += SPS) st.advance();
            }
            // exact path: drain after every unit; fast path: drain after this pipeline's last unit of the tile run
            const bool last_owned = (w.u + 2 > w.run_last());
            if (!fast || last_owned) {
                if (ptx::elect_one()) ptx::umma_commit(ptx::smem_u32(&s_accfull[d][ab]));
                __syncwarp();
                ++ndrain;
                open = false;
            } else {
                open = true;
            }
        }
    } else if (warp >= 4 && warp < 12) {
        // ===================== decode set d: packed 2-bit codes -> u8 A operand in TMEM =====================
        const int d = (warp - 4) >> 2;
        const int dw = (warp - 4) & 3;
        const int m = dw * 32 + lane;  // weight row inside the strip == TMEM lane
        Walk w;
        w.init(p, u0, u1);
        Ring st;
        st.init(S);
        uint32_t a_ph = 0;
        const uint32_t taddr0 = tmem_base + ((uint32_t) (dw * 32) << 16) + A_COL0 + d * A_COLS;
        for (; w.valid(); w.next()) {
            if (w.owner() != d) continue;
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                bool a_free = false;
                for (int s0 = 0; s0 < na; s0 += SPS) {
                    const int ns = min(SPS, na - s0);
                    ptx::mbar_wait(ptx::smem_u32(&s_full[d][st.i]), st.ph);
                    const uint32_t w_s = smem_base + (d * S + st.i) * p.stage_bytes + SPS * ACT_BYTES + m * 16;
                    uint4 cw[2 * SPS];
#pragma unroll
                    for (int c = 0; c < 2 * SPS; c++)
                        if (c < 2 * ns) cw[c] = ptx::lds128(w_s + c * TILE_BYTES);
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[d][st.i]));  // weight part of the stage consumed
                    st.advance();
                    if (!a_free) {
                        ptx::mbar_wait(ptx::smem_u32(&s_aempty[d]), a_ph ^ 1);
                        ptx::tc_fence_after();
                        a_free = true;
                    }
#pragma unroll
                    for (int c = 0; c < 2 * SPS; c++) {
                        if (c < 2 * ns) {
                            const uint32_t wv[4] = {cw[c].x, cw[c].y, cw[c].z, cw[c].w};
                            uint32_t r[16];
                            if (p.ablate & 1) {
#pragma unroll
                                for (int j = 0; j < 16; j++) r[j] = wv[j & 3];
                            } else {
#pragma unroll
                                for (int j = 0; j < 4; j++) {
                                    // shifts as multiply-high on the FMA pipe, masks on the ALU pipe: both pipes work
                                    r[4 * j + 0] = wv[j] & 0x03030303u;
                                    r[4 * j + 1] = __umulhi(wv[j], 1u << 30) & 0x03030303u;
                                    r[4 * j + 2] = __umulhi(wv[j], 1u << 28) & 0x03030303u;
                                    r[4 * j + 3] = __umulhi(wv[j], 1u << 26) & 0x03030303u;
                                }
                            }
                            if (!(p.ablate & 2)) ptx::tmem_st_32x32b_x16(taddr0 + s0 * 32 + c * 16, r);
                        }
                    }
                }
                ptx::tmem_st_wait();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_afull[d]));
                a_ph ^= 1;
            }
        }
    } else if (warp < 4) {
        // ===================== epilogue =====================
        const int ew = warp;
        const int m = ew * 32 + lane;
        const int et = threadIdx.x;  // 0..127
        // exact-path running totals: shared memory when they fit next to the stages, else this CTA's third workspace slot
        int32_t *tots = p.tots_off >= 0
                            ? reinterpret_cast<int32_t *>(smem_raw + (smem_base - ptx::smem_u32(smem_raw)) + p.tots_off)
                            : p.g.part + ((int64_t) gridDim.x * 2 + blockIdx.x) * (NT * 128);
        uint32_t ndrain[2] = {0, 0};
        Walk w;
        w.init(p, u0, u1);
        while (w.valid()) {
            // ---- one tile run: units [ra, rb] of one output tile ----
            const int ra = w.u, rb = w.run_last();
            const int g_begin = w.g, g_end = w.g + (rb - ra);
            const bool fast = tile_fast(w.tt);
            const bool complete = (g_begin == 0) && (g_end == p.G - 1);
            const bool run_is_first = (ra == u0);
            const int64_t n = (int64_t) w.rt * TILE_ROWS + m;
            const int64_t t0 = (int64_t) w.tt * NT;
            const int tile = w.tile;
            const int32_t *pre = p.g.act.gpre + (t0 + et) * (p.G + 1);
            int32_t *slot = p.g.part + ((int64_t) blockIdx.x * 2 + (run_is_first ? 0 : 1)) * (NT * 128);
            if (et < NT) s_sc[et] = __ldg(p.g.act.scales + t0 + et);

            // final values of 16 columns: store (complete tile) or park in the workspace slot (split tile)
            auto emit = [&](int c0, const int32_t (&vals)[16]) {
                if (complete) {
                    if (n < p.g.N) {
#pragma unroll
                        for (int i = 0; i < 16; i++) {
                            if (t0 + c0 + i < p.g.T) {
                                float v = (float) vals[i] * s_sc[c0 + i];
                                if (p.g.w_scale != 1.0f) v *= p.g.w_scale;
                                store_out(p.g, (t0 + c0 + i) * p.g.ldo + n, v);
                            }
                        }
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; i++) __stcg(slot + (c0 + i) * 128 + m, vals[i]);
                }
            };

            if (fast) {
                // single drain: both pipelines' accumulators are summed as they are read
                if (et < NT) s_gs[et] = __ldg(pre + g_end + 1) - __ldg(pre + g_begin);
                ptx::bar_sync(1, 128);
                const bool has1 = rb > ra;  // pipeline of unit ra always takes part; the other one iff the run has >= 2 units
                const int d0 = (ra - u0) & 1, d1 = d0 ^ 1;
                const uint32_t b0 = ndrain[d0] % NACC, b1 = ndrain[d1] % NACC;
                ptx::mbar_wait(ptx::smem_u32(&s_accfull[d0][b0]), (ndrain[d0] / NACC) & 1);
                if (has1) ptx::mbar_wait(ptx::smem_u32(&s_accfull[d1][b1]), (ndrain[d1] / NACC) & 1);
                ptx::tc_fence_after();
                const uint32_t ta0 = tmem_base + ((uint32_t) (ew * 32) << 16) + (d0 * NACC + b0) * NT;
                const uint32_t ta1 = tmem_base + ((uint32_t) (ew * 32) << 16) + (d1 * NACC + b1) * NT;
#pragma unroll 1
                for (int c0 = 0; c0 < NT; c0 += 16) {
                    uint32_t v0[16], v1[16];
                    ptx::tmem_ld_32x32b_x16(ta0 + c0, v0);
                    if (has1) ptx::tmem_ld_32x32b_x16(ta1 + c0, v1);
                    ptx::tmem_ld_wait();
                    int32_t vals[16];
#pragma unroll
                    for (int i = 0; i < 16; i++)
                        vals[i] = (int32_t) v0[i] + (has1 ? (int32_t) v1[i] : 0) - s_gs[c0 + i];
                    emit(c0, vals);
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    ptx::mbar_arrive(ptx::smem_u32(&s_accempty[d0][b0]));
                    if (has1) ptx::mbar_arrive(ptx::smem_u32(&s_accempty[d1][b1]));
                }
                ++ndrain[d0];
                if (has1) ++ndrain[d1];
                for (int u = ra; u <= rb; u++) w.next();
            } else {
                // exact path: one drain per unit (K-group), int16 wrap, running totals in shared memory
                for (int u = ra; u <= rb; u++, w.next()) {
                    const int d = (u - u0) & 1;
                    const uint32_t b = ndrain[d] % NACC;
                    if (et < NT) s_gs[et] = __ldg(pre + w.g + 1) - __ldg(pre + w.g);
                    ptx::bar_sync(1, 128);
                    ptx::mbar_wait(ptx::smem_u32(&s_accfull[d][b]), (ndrain[d] / NACC) & 1);
                    ptx::tc_fence_after();
                    const uint32_t ta = tmem_base + ((uint32_t) (ew * 32) << 16) + (d * NACC + b) * NT;
#pragma unroll 1
                    for (int c0 = 0; c0 < NT; c0 += 16) {
                        uint32_t v[16];
                        ptx::tmem_ld_32x32b_x16(ta + c0, v);
                        ptx::tmem_ld_wait();
                        int32_t vals[16];
#pragma unroll
                        for (int i = 0; i < 16; i++) {
                            const int32_t x = wrap16((int32_t) v[i] - s_gs[c0 + i]);
                            vals[i] = (u == ra) ? x : tots[(c0 + i) * 128 + et] + x;
                        }
                        if (u == rb) {
                            emit(c0, vals);
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; i++) tots[(c0 + i) * 128 + et] = vals[i];
                        }
                    }
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_accempty[d][b]));
                    ++ndrain[d];
                    ptx::bar_sync(1, 128);  // s_gs is rewritten by the next unit
                }
            }

            if (!complete) {
                // split tile (stream-K): partial parked above; the last CTA to arrive sums all slots and finishes
                __threadfence();
                ptx::bar_sync(1, 128);
                if (et == 0) {
                    const int ndone = g_end - g_begin + 1;
                    const int old = atomicAdd(p.g.counters + tile, ndone);
                    s_last = (old + ndone == p.G);
                }
                ptx::bar_sync(1, 128);
                if (s_last) {
                    __threadfence();
                    const long long ua = (long long) tile * p.G, ub = ua + p.G - 1;
                    const int b_lo = (int) (((ua + 1) * gridDim.x - 1) / p.U);
                    const int b_hi = (int) (((ub + 1) * gridDim.x - 1) / p.U);
#pragma unroll 1
                    for (int c0 = 0; c0 < NT; c0 += 16) {
                        int32_t acc[16];
#pragma unroll
                        for (int i = 0; i < 16; i++) acc[i] = 0;
                        for (int b = b_lo; b <= b_hi; b++) {
                            const int b_u0 = (int) (((long long) b * p.U) / gridDim.x);
                            const int b_slot = (b_u0 / p.G == tile) ? 0 : 1;
                            const int32_t *ps = p.g.part + ((int64_t) b * 2 + b_slot) * (NT * 128) + m;
                            int32_t a[16];
#pragma unroll
                            for (int i = 0; i < 16; i++) a[i] = __ldcg(ps + (c0 + i) * 128);
#pragma unroll
                            for (int i = 0; i < 16; i++) acc[i] += a[i];
                        }
                        if (n < p.g.N) {
#pragma unroll
                            for (int i = 0; i < 16; i++) {
                                if (t0 + c0 + i < p.g.T) {
                                    float v = (float) acc[i] * s_sc[c0 + i];
                                    if (p.g.w_scale != 1.0f) v *= p.g.w_scale;
                                    store_out(p.g, (t0 + c0 + i) * p.g.ldo + n, v);
                                }
                            }
                        }
                    }
                    if (et == 0) p.g.counters[tile] = 0;
                }
            }
            ptx::bar_sync(1, 128);  // s_sc / s_gs / s_last are reused by the next run
        }
    }

    // ---- teardown ----
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 15) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 512);
    }
}

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}

template <int NT, int SPS>
static cudaError_t launch_one(Params &p, int grid, cudaStream_t s) {
    const int budget = env_int("VLUT_TC_SMEM_KB", 200) * 1024;
    p.stage_bytes = SPS * (NT * 128 + SUB_W_BYTES);  // multiple of 1024
    // exact-path running totals: in shared memory up to NT = 64, in the global workspace for NT = 128
    const int tots_bytes = (NT <= 64) ? NT * 128 * 4 : 0;
    int ns = (budget - tots_bytes - 1024) / p.stage_bytes;
    if (ns > MAX_STAGES) ns = MAX_STAGES;
    int forced = 2 * env_int("VLUT_TC_STAGES", 0);  // per pipeline
    if (forced > 0 && forced < 2 * SPA / SPS) forced = 2 * SPA / SPS;
    if (forced > 0 && forced < ns) ns = forced;
    ns &= ~1;  // two equal rings
    if (ns < 2 * SPA / SPS) return cudaErrorInvalidValue;  // every stage of one A block is live at once, per pipeline
    p.nstages = ns;
    p.tots_off = tots_bytes ? ns * p.stage_bytes : -1;
    const size_t smem = (size_t) ns * p.stage_bytes + tots_bytes + 1024;
    auto kern = kernel<NT, SPS>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess) return e;
    kern<<<grid, THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace tc2

bool gemm_tc2_supported(int type, int64_t K, int64_t N, int64_t T) {
    return type_valid(type) && T >= 1 && K >= 4 && N >= 1 && ceil_div(K, group_size(type)) <= 64;
}

cudaError_t launch_gemm_tc2(const GemmArgs &g, cudaStream_t s, int *launches) {
    tc2::Params p{};
    p.g = g;
    p.KB = (int) (g.act.Kpad / 128);
    p.gsub = g.act.gsize / 128;
    p.G = g.act.G;
    p.RT = (int) (g.Npad / TILE_ROWS);
    const int NT = g.act.NT;
    p.TT = (int) (g.act.Tpad / NT);
    p.U = p.RT * p.TT * p.G;
    p.allow_fast = tc2::env_int("VLUT_TC_FAST", 1);
    p.ablate = tc2::env_int("VLUT_TC_ABLATE", 0);
    int grid = g.sm_count;
    const int tiles = p.RT * p.TT;
    if (tiles <= g.sm_count && tiles * 10 >= g.sm_count * 7) grid = tiles;  // whole tiles per CTA: no split-tile fix-up
    grid = tc2::env_int("VLUT_TC_GRID", grid);
    if (grid > p.U) grid = p.U;
    if (grid < 1) grid = 1;
    if (launches) *launches += 1;
    switch (NT) {
        case 16: return tc2::launch_one<16, 4>(p, grid, s);
        case 32: return tc2::launch_one<32, 4>(p, grid, s);
        case 64: return tc2::launch_one<64, 2>(p, grid, s);
        case 128: return tc2::launch_one<128, 2>(p, grid, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace vlut
