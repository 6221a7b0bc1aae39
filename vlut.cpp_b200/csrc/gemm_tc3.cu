// gemm_tc3.cu -- VLUT_ALGO_TC: the Vec-LUT GEMM on tcgen05 (kind::i8): 256 weight rows per CTA, two decode+issue
// pipelines that SHARE every activation stage.
//
// Same algorithm, layouts and exactness contract as gemm_tc.cu.  What the measurements of the
// earlier kernels showed (profiles/r1_microbench_b200.txt): with every arithmetic instruction removed the kernel was
// only ~15 % faster -- time was in the load / synchronisation skeleton.  At large T a 128 x 128 x 512 unit needs 80 KB
// of shared-memory fill, 64 KB of it activations that every 128-row tile re-reads: 78 B/clk/SM at full tensor rate
// against ~42 B/clk/SM of L2->SM bandwidth, and only one A block of stages in flight.  Here
//   * a CTA works on a PAIR of row tiles (256 weight rows) against one token tile: pipeline d (decode set d, issuer d,
//     its own TMEM A stage and accumulators) owns row tile 2*rp + d; both pipelines read the SAME activation stage, so
//     activation fill per MAC halves (a unit = 32 KB weights + NT x 512 B activations per K-group);
//   * shared-memory stages are 256 k (NT = 64/128) deep while the decode -> MMA hand-off stays at 512 k: two A blocks of
//     stages in flight;
//   * programmatic dependent launch: the kernel lets its successor start early and prefetches its first weight stages
//     BEFORE waiting for its predecessor (the activation quantiser, or the previous GEMM): static weights need no
//     ordering, only activations, the epilogue's inputs and all global writes do.
//
//   warps 0-3   epilogue              warps 4-7  decode set 0     warps 8-11  decode set 1
//   warp 12     TMA producer          warp 13    MMA issuer 0     warp 14     MMA issuer 1      warp 15  TMEM allocator
#include <stdlib.h>

#include "ptx.cuh"
#include "vlut_common.cuh"

namespace vlut {

namespace tc3 {

constexpr int THREADS = 512;
constexpr int SPA = 4;             // 128-k sub-blocks per A stage / decode->MMA hand-off
constexpr int A_COLS = SPA * 32;   // TMEM columns of one A stage
constexpr int MAX_STAGES = 8;
constexpr int SUB_W_BYTES = 4096;

struct Params {
    GemmArgs g;
    int KB, G, gsub, RT, RP, TT, U;
    int nstages, stage_bytes;
    int allow_fast, pdl;
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

// units of this CTA: unit = (row-tile pair rp, token tile tt, K-group g), ordered [rp][tt][g]
struct Walk {
    int u, u0, u1, G, TT, gsub, KB;
    int tp, g, rp, tt, kb0, nsub;
    __device__ __forceinline__ void init(const Params &p, int u0_, int u1_) {
        u = u0 = u0_; u1 = u1_; G = p.G; TT = p.TT; gsub = p.gsub; KB = p.KB;
        tp = u / G; g = u - tp * G;
        rp = tp / TT; tt = tp - rp * TT;
        refresh();
    }
    __device__ __forceinline__ void refresh() { kb0 = g * gsub; nsub = min(gsub, KB - kb0); }
    __device__ __forceinline__ bool valid() const { return u < u1; }
    __device__ __forceinline__ int run_last() const { return min(u1 - 1, u - g + G - 1); }
    __device__ __forceinline__ void next() {
        ++u;
        if (++g == G) { g = 0; ++tp; if (++tt == TT) { tt = 0; ++rp; } }
        refresh();
    }
};

struct Ring {
    int i, n;
    uint32_t ph;
    __device__ __forceinline__ void init(int n_) { i = 0; n = n_; ph = 0; }
    __device__ __forceinline__ void advance() { if (++i == n) { i = 0; ph ^= 1; } }
};

// OUT: epilogue store flavour (0 local, 1 peer scatter, 2 NVSwitch multicast)
template <int NT, int SPS, int OUT>
__global__ void __launch_bounds__(THREADS, 1) kernel(const __grid_constant__ Params p) {
    static_assert(SPA % SPS == 0, "an A stage is a whole number of shared-memory stages");
    constexpr int NACC = (NT == 128) ? 1 : 2;  // accumulator buffers per pipeline
    constexpr int ACT_BYTES = NT * 128;
    constexpr int W_OFF = SPS * ACT_BYTES;          // stage: [activations][weights pipeline 0][weights pipeline 1]
    constexpr int W_PIPE = SPS * SUB_W_BYTES;
    constexpr uint32_t A_COL0 = 2 * NACC * NT;
    static_assert(A_COL0 + 2 * A_COLS <= 512, "TMEM budget");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_full[MAX_STAGES], s_empty[MAX_STAGES];
    __shared__ __align__(8) uint64_t s_afull[2], s_aempty[2];
    __shared__ __align__(8) uint64_t s_accfull[2][2], s_accempty[2][2];
    __shared__ uint32_t s_tmem_base;
    __shared__ int s_last;
    __shared__ int32_t s_gs[NT];
    __shared__ float s_sc[NT];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = p.nstages;
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;

    // let the next kernel of the stream start its own prologue / weight prefetch as early as possible
    if (p.pdl) ptx::pdl_launch_dependents();

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) {
            ptx::mbar_init(ptx::smem_u32(&s_full[s]), 1);
            ptx::mbar_init(ptx::smem_u32(&s_empty[s]), 8 + 2);  // 2 x 4 decode warps + 2 issuers
        }
        for (int d = 0; d < 2; d++) {
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
    float *const out0 = (OUT == 0) ? p.g.out : p.g.outs[0];

    const int u0 = (int) (((int64_t) blockIdx.x * p.U) / gridDim.x);
    const int u1 = (int) (((int64_t) (blockIdx.x + 1) * p.U) / gridDim.x);
    auto tile_fast = [&](int tt) -> bool { return p.allow_fast && (__ldg(p.g.act.tile_unsafe + tt) != p.g.act.epoch); };
    auto pipe_valid = [&](int rp, int d) -> bool { return 2 * rp + d < p.RT; };

    if (warp == 12) {
        // ===================== TMA producer =====================
        const uint64_t pol_w = ptx::policy_evict_first();
        const uint64_t pol_a = ptx::policy_evict_last();
        Walk w;
        w.init(p, u0, u1);
        Ring st;
        st.init(S);
        // With PDL the first S stages get their WEIGHTS before griddepcontrol.wait (weights are static), their
        // activations (written by the predecessor kernel) after it.
        int early = p.pdl ? S : 0;
        bool waited = !p.pdl;
        Walk wa = w;  // activation issue trails the weight issue while `early` stages are outstanding
        Ring sta = st;
        auto issue = [&](Walk &ww, Ring &rr, bool do_w, bool do_a, int ab0, int s0, int ns) {
            const bool v1 = pipe_valid(ww.rp, 1);
            const uint32_t full = ptx::smem_u32(&s_full[rr.i]);
            const uint32_t dst = smem_base + rr.i * p.stage_bytes;
            if (do_w) {
                ptx::mbar_arrive_expect_tx(full, ns * (ACT_BYTES + SUB_W_BYTES * (v1 ? 2 : 1)));
                const uint8_t *w0 = p.g.w + ((int64_t) (2 * ww.rp) * p.KB + ww.kb0 + ab0 + s0) * SUB_W_BYTES;
                ptx::bulk_g2s_hint(dst + W_OFF, w0, ns * SUB_W_BYTES, full, pol_w);
                if (v1) ptx::bulk_g2s_hint(dst + W_OFF + W_PIPE, w0 + (int64_t) p.KB * SUB_W_BYTES, ns * SUB_W_BYTES, full, pol_w);
            }
            if (do_a) {
                const int8_t *a = p.g.act.q_sw + ((int64_t) ww.tt * p.KB + ww.kb0 + ab0 + s0) * ACT_BYTES;
                ptx::bulk_g2s_hint(dst, a, ns * ACT_BYTES, full, pol_a);
            }
        };
        for (; w.valid(); w.next()) {
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                for (int s0 = 0; s0 < na; s0 += SPS) {
                    const int ns = min(SPS, na - s0);
                    if (early > 0) {
                        if (ptx::elect_one()) issue(w, st, true, false, ab0, s0, ns);  // fresh barrier: no empty wait needed
                        __syncwarp();
                        st.advance();
                        --early;
                        continue;
                    }
                    if (!waited) {
                        // all early weight loads are in flight: now order against the predecessor and send activations
                        ptx::pdl_wait();
                        waited = true;
                        Walk wb = wa;
                        Ring rb = sta;
                        for (int e = 0; e < S && wb.valid();) {
                            for (int b0 = 0; b0 < wb.nsub && e < S; b0 += SPA) {
                                const int nb = min(SPA, wb.nsub - b0);
                                for (int t0 = 0; t0 < nb && e < S; t0 += SPS, ++e) {
                                    if (ptx::elect_one()) issue(wb, rb, false, true, b0, t0, min(SPS, nb - t0));
                                    __syncwarp();
                                    rb.advance();
                                }
                            }
                            wb.next();
                        }
                    }
                    ptx::mbar_wait(ptx::smem_u32(&s_empty[st.i]), st.ph ^ 1);
                    if (ptx::elect_one()) issue(w, st, true, true, ab0, s0, ns);
                    __syncwarp();
                    st.advance();
                }
            }
        }
        if (!waited) {
            // fewer than S stages in total: activations still have to be sent
            ptx::pdl_wait();
            Walk wb = wa;
            Ring rb = sta;
            for (; wb.valid(); wb.next())
                for (int b0 = 0; b0 < wb.nsub; b0 += SPA) {
                    const int nb = min(SPA, wb.nsub - b0);
                    for (int t0 = 0; t0 < nb; t0 += SPS) {
                        if (ptx::elect_one()) issue(wb, rb, false, true, b0, t0, min(SPS, nb - t0));
                        __syncwarp();
                        rb.advance();
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
        uint32_t a_ph = 0;    // phase of this pipeline's A stage
        uint32_t ndrain = 0;  // drains committed by this pipeline so far
        if (p.pdl) ptx::pdl_wait();  // tile_fast below is written by the predecessor (quantise) kernel
        int run_tp = -1;
        bool fast = false, open = false;
        for (; w.valid(); w.next()) {
            const bool valid = pipe_valid(w.rp, d);
            if (w.tp != run_tp) {
                run_tp = w.tp;
                fast = tile_fast(w.tt);
            }
            const uint32_t ab = ndrain % NACC;
            if (valid && !open) {
                ptx::mbar_wait(ptx::smem_u32(&s_accempty[d][ab]), ((ndrain / NACC) & 1) ^ 1);
                ptx::tc_fence_after();
            }
            const uint32_t d_tmem = tmem_base + (d * NACC + ab) * NT;
            const uint32_t a_tmem = tmem_base + A_COL0 + d * A_COLS;
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                if (!valid) {
                    // odd row-tile count: this pipeline idles, but the shared stages still need its release
                    for (int s0 = 0; s0 < na; s0 += SPS) {
                        ptx::mbar_wait(ptx::smem_u32(&s_full[st.i]), st.ph);
                        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[st.i]));
                        __syncwarp();
                        st.advance();
                    }
                    continue;
                }
                ptx::mbar_wait(ptx::smem_u32(&s_afull[d]), a_ph);
                a_ph ^= 1;
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
                                ptx::umma_i8_ts(d_tmem, a_tmem + (s0 + sb) * 32 + ks * 8, b_desc + (uint64_t) (ks * 2), idesc,
                                                accumulate);
                            }
                        }
                        ptx::umma_commit(ptx::smem_u32(&s_empty[st2.i]));
                        st2.advance();
                    }
                    ptx::umma_commit(ptx::smem_u32(&s_aempty[d]));
                }
                __syncwarp();
                for (int s0 = 0; s0 < na; s0 += SPS) st.advance();
            }
            if (!valid) continue;
            // exact path: drain after every unit; fast path: drain after the last unit of the tile run
            if (!fast || w.u == w.run_last()) {
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
            const bool valid = pipe_valid(w.rp, d);
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                bool a_free = false;
                for (int s0 = 0; s0 < na; s0 += SPS) {
                    const int ns = min(SPS, na - s0);
                    ptx::mbar_wait(ptx::smem_u32(&s_full[st.i]), st.ph);
                    const uint32_t w_s = smem_base + st.i * p.stage_bytes + W_OFF + d * W_PIPE + m * 16;
                    uint4 cw[2 * SPS];
                    if (valid) {
#pragma unroll
                        for (int c = 0; c < 2 * SPS; c++)
                            if (c < 2 * ns) cw[c] = ptx::lds128(w_s + c * TILE_BYTES);
                    }
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[st.i]));  // this set's weights are consumed
                    st.advance();
                    if (!valid) continue;
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
#pragma unroll
                            for (int j = 0; j < 4; j++) {
                                // shifts as multiply-high on the FMA pipe, masks on the ALU pipe: both pipes work
                                r[4 * j + 0] = wv[j] & 0x03030303u;
                                r[4 * j + 1] = __umulhi(wv[j], 1u << 30) & 0x03030303u;
                                r[4 * j + 2] = __umulhi(wv[j], 1u << 28) & 0x03030303u;
                                r[4 * j + 3] = __umulhi(wv[j], 1u << 26) & 0x03030303u;
                            }
                            ptx::tmem_st_32x32b_x16(taddr0 + s0 * 32 + c * 16, r);
                        }
                    }
                }
                if (!valid) continue;
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
        // everything the epilogue reads from global memory was written by predecessor kernels; it also owns all
        // global writes of this kernel
        if (p.pdl) ptx::pdl_wait();
        uint32_t ndrain[2] = {0, 0};
        Walk w;
        w.init(p, u0, u1);
        while (w.valid()) {
            // ---- one tile run: units [ra, rb] of one pair of output tiles ----
            const int ra = w.u, rb = w.run_last();
            const int g_begin = w.g, g_end = w.g + (rb - ra);
            const bool fast = tile_fast(w.tt);
            const bool complete = (g_begin == 0) && (g_end == p.G - 1);
            const int run_slot = (ra == u0) ? 0 : 1;
            const int rp = w.rp, tp = w.tp;
            const int64_t t0 = (int64_t) w.tt * NT;
            const int tt = w.tt;
            const int32_t *pre = p.g.act.gpre + (t0 + et) * (p.G + 1);
            if (et < NT) s_sc[et] = __ldg(p.g.act.scales + t0 + et);
            const int nd = pipe_valid(rp, 1) ? 2 : 1;

            // final values of 16 columns of pipeline d's tile: store (complete tile) or park in the workspace slot
            auto emit = [&](int d, int c0, const int32_t (&vals)[16]) {
                const int64_t n = (int64_t) (2 * rp + d) * TILE_ROWS + m;
                if (complete) {
                    if (n < p.g.N) {
#pragma unroll
                        for (int i = 0; i < 16; i++) {
                            if (t0 + c0 + i < p.g.T) {
                                float v = (float) vals[i] * s_sc[c0 + i];
                                if (p.g.w_scale != 1.0f) v *= p.g.w_scale;
                                store_out<OUT>(out0, p.g.outs, p.g.n_out, (t0 + c0 + i) * p.g.ldo + n, v);
                            }
                        }
                    }
                } else {
                    int32_t *slot = p.g.part + ((int64_t) blockIdx.x * 4 + run_slot * 2 + d) * (NT * 128);
#pragma unroll
                    for (int i = 0; i < 16; i++) __stcg(slot + (c0 + i) * 128 + m, vals[i]);
                }
            };

            if (fast) {
                if (et < NT) s_gs[et] = __ldg(pre + g_end + 1) - __ldg(pre + g_begin);
                ptx::bar_sync(1, 128);
                for (int d = 0; d < nd; d++) {
                    const uint32_t b = ndrain[d] % NACC;
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
                        for (int i = 0; i < 16; i++) vals[i] = (int32_t) v[i] - s_gs[c0 + i];
                        emit(d, c0, vals);
                    }
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_accempty[d][b]));
                    ++ndrain[d];
                }
                for (int u = ra; u <= rb; u++) w.next();
            } else {
                // exact path: one drain per unit (K-group), int16 wrap, running totals in this CTA's global scratch
                for (int u = ra; u <= rb; u++, w.next()) {
                    if (et < NT) s_gs[et] = __ldg(pre + w.g + 1) - __ldg(pre + w.g);
                    ptx::bar_sync(1, 128);
                    for (int d = 0; d < nd; d++) {
                        int32_t *tots = p.g.part + ((int64_t) gridDim.x * 4 + (int64_t) blockIdx.x * 2 + d) * (NT * 128);
                        const uint32_t b = ndrain[d] % NACC;
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
                                emit(d, c0, vals);
                            } else {
#pragma unroll
                                for (int i = 0; i < 16; i++) tots[(c0 + i) * 128 + et] = vals[i];
                            }
                        }
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_accempty[d][b]));
                        ++ndrain[d];
                    }
                    ptx::bar_sync(1, 128);  // s_gs is rewritten by the next unit
                }
            }

            if (!complete) {
                // split tiles (stream-K): partials parked above; per tile, the last CTA to arrive sums all slots
                __threadfence();
                ptx::bar_sync(1, 128);
                for (int d = 0; d < nd; d++) {
                    const int tile = (2 * rp + d) * p.TT + tt;
                    if (et == 0) {
                        const int ndone = g_end - g_begin + 1;
                        const int old = atomicAdd(p.g.counters + tile, ndone);
                        s_last = (old + ndone == p.G);
                    }
                    ptx::bar_sync(1, 128);
                    if (s_last) {
                        __threadfence();
                        const int64_t n = (int64_t) (2 * rp + d) * TILE_ROWS + m;
                        const long long ua = (long long) tp * p.G, ub = ua + p.G - 1;
                        const int b_lo = (int) (((ua + 1) * gridDim.x - 1) / p.U);
                        const int b_hi = (int) (((ub + 1) * gridDim.x - 1) / p.U);
#pragma unroll 1
                        for (int c0 = 0; c0 < NT; c0 += 16) {
                            int32_t acc[16];
#pragma unroll
                            for (int i = 0; i < 16; i++) acc[i] = 0;
                            for (int b = b_lo; b <= b_hi; b++) {
                                const int b_u0 = (int) (((long long) b * p.U) / gridDim.x);
                                const int b_slot = (b_u0 / p.G == tp) ? 0 : 1;
                                const int32_t *ps = p.g.part + ((int64_t) b * 4 + b_slot * 2 + d) * (NT * 128) + m;
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
                                        store_out<OUT>(out0, p.g.outs, p.g.n_out, (t0 + c0 + i) * p.g.ldo + n, v);
                                    }
                                }
                            }
                        }
                        if (et == 0) p.g.counters[tile] = 0;
                    }
                    ptx::bar_sync(1, 128);  // s_last is rewritten for the second tile of the pair
                }
            }
            ptx::bar_sync(1, 128);  // s_sc / s_gs are reused by the next run
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

template <int NT, int SPS, int OUT>
static cudaError_t launch_out(Params &p, int grid, cudaStream_t s) {
    const int budget = env_int("VLUT_TC_SMEM_KB", 200) * 1024;
    p.stage_bytes = SPS * (NT * 128 + 2 * SUB_W_BYTES);  // multiple of 1024
    int ns = (budget - 1024) / p.stage_bytes;
    if (ns > MAX_STAGES) ns = MAX_STAGES;
    int forced = env_int("VLUT_TC_STAGES", 0);
    if (forced > 0 && forced < SPA / SPS) forced = SPA / SPS;
    if (forced > 0 && forced < ns) ns = forced;
    if (ns < SPA / SPS) return cudaErrorInvalidValue;  // every stage of one A block is live at once
    p.nstages = ns;
    const size_t smem = (size_t) ns * p.stage_bytes + 1024;
    auto kern = kernel<NT, SPS, OUT>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) grid);
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = p.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, p);
}

template <int NT, int SPS>
static cudaError_t launch_one(Params &p, int grid, cudaStream_t s) {
    if (p.g.n_out == 0) return launch_out<NT, SPS, 0>(p, grid, s);
    return p.g.multicast ? launch_out<NT, SPS, 2>(p, grid, s) : launch_out<NT, SPS, 1>(p, grid, s);
}

}  // namespace tc3

bool gemm_tc3_supported(int type, int64_t K, int64_t N, int64_t T) {
    return type_valid(type) && T >= 1 && K >= 4 && N >= 1 && ceil_div(K, group_size(type)) <= 64;
}

cudaError_t launch_gemm_tc3(const GemmArgs &g, cudaStream_t s, int *launches) {
    if (g.act.sw_tile != g.act.NT) return cudaErrorInvalidValue;  // wants full-height swizzled activation tiles
    tc3::Params p{};
    p.g = g;
    p.KB = (int) (g.act.Kpad / 128);
    p.gsub = g.act.gsize / 128;
    p.G = g.act.G;
    p.RT = (int) (g.Npad / TILE_ROWS);
    p.RP = (p.RT + 1) / 2;
    const int NT = g.act.NT;
    p.TT = (int) (g.act.Tpad / NT);
    p.U = p.RP * p.TT * p.G;
    p.allow_fast = tc3::env_int("VLUT_TC_FAST", 1);
    p.pdl = tc3::env_int("VLUT_PDL", 1);
    int grid = g.sm_count;
    const int pairs = p.RP * p.TT;
    if (pairs <= g.sm_count && pairs * 10 >= g.sm_count * 7) grid = pairs;  // whole tile pairs per CTA: no fix-up
    grid = tc3::env_int("VLUT_TC_GRID", grid);
    if (grid > p.U) grid = p.U;
    if (grid < 1) grid = 1;
    if (launches) *launches += 1;
    switch (NT) {
        case 16: return tc3::launch_one<16, 4>(p, grid, s);
        case 32: return tc3::launch_one<32, 4>(p, grid, s);
        case 64: return tc3::launch_one<64, 2>(p, grid, s);
        case 128: return tc3::launch_one<128, 2>(p, grid, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace vlut
