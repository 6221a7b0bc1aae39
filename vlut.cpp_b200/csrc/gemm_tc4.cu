// gemm_tc4.cu -- VLUT_ALGO_TC: the production Vec-LUT GEMM on tcgen05 (kind::i8) for sm_100a.
//
// Algorithm, layouts and exactness contract: DESIGN.md section 4.1 (and gemm_tc.cu, the first-generation kernel kept as
// VLUT_ALGO_TC1).  This generation is shaped by the in-kernel timelines of the first one (scripts/debug_counters.py,
// profiles/r1_timeline_tc1.txt): at every token count the DECODE stage was the bottleneck -- ~1300-1750 clk per 512-k A
// block for 4 warps (232 shift/mask instructions per warp on the ALU pipe at 2 clk each, 8 tcgen05.st, LDS, barrier
// round trips) against 288 clk (N=16) ... 1024 clk (N=128) of tensor-pipe time and ~700 clk of HBM time per block.
//   * EIGHT decode warps share every A block: warps (q, h) -- q the TMEM lane quadrant, h the half -- unpack the two
//     halves of each shared-memory stage side by side into disjoint TMEM columns of the same A stage;
//   * shifts are issued as multiply-high (IMAD.HI on the FMA pipe), masks as LOP3 (ALU pipe): both pipes work;
//   * the epilogue is register-light (16 accumulator columns at a time, no per-thread running totals: the exact path
//     keeps its per-K-group int16-wrap totals in an L2-resident scratch tile), so 480 threads fit;
//   * programmatic dependent launch: successors start early; the first stages' WEIGHTS are requested before
//     griddepcontrol.wait (static data), activations / epilogue inputs / all global writes after it;
//   * the epilogue store flavour (local, NVLink peer scatter, NVSwitch multicast) is a template parameter.
//
//   warps 0-3   epilogue            warps 4-11  decode (quadrant q = w%4, half h = (w-4)/4)
//   warp 12     activation producer warp 13     MMA issuer          warp 14  TMEM allocator, then weight producer
#include <stdlib.h>

#include "epilogue.cuh"
#include "ptx.cuh"
#include "vlut_common.cuh"

namespace vlut {

long long *debug_buffer(cudaStream_t s, int skip);  // gemm_tc.cu

namespace tc4 {

__host__ __device__ constexpr int npw_for(int dg) { return dg == 3 ? 3 : 2; }  // weight-producer warps
// 4 epilogue + decode + activation producer + MMA + weight producers (the first one also allocates TMEM)
__host__ __device__ constexpr int threads_for(int dg) { return (6 + (dg == 3 ? 12 : 8) + npw_for(dg)) * 32; }
constexpr int SPA = 4;             // 128-k sub-blocks per A stage / decode->MMA hand-off
constexpr int A_COLS = SPA * 32;   // TMEM columns of one A stage
constexpr int MAX_STAGES = 10;
constexpr int SUB_W_BYTES = 4096;

struct Params {
    GemmArgs g;
    int KB, G, gsub, RT, TT, U;
    int nstages, stage_bytes;
    int allow_fast, pdl;
    int whole;  // 1: every CTA gets whole output tiles (no fix-up); 0: stream-K over (tile, K-group) units
    long long *dbg;  // optional timeline buffer (VLUT_TC_DEBUG), else null
    int dbg_skip;    // events to skip per role before recording
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

// units of this CTA: unit = (row tile rt, token tile tt, K-group g), ordered [rt][tt][g]
struct Walk {
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
    __device__ __forceinline__ int run_last() const { return min(u1 - 1, u - g + G - 1); }
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

// OUT: epilogue store flavour (0 local, 1 peer scatter, 2 NVSwitch multicast)
// DG: decode organisation.  1: eight warps share every A block (two A stages).  3 (small token tiles): three groups of four
// warps take every third A block, one whole block (= one shared-memory stage) per group and turn, three A stages: three
// decode chains (stage wait, LDS, unpack, tcgen05.st, wait::st, hand-over: ~1600 clk of mostly latency) run side by side.
template <int NT, int SPS, int OUT, int DG>
__global__ void __launch_bounds__(threads_for(DG), 1) kernel(const __grid_constant__ Params p) {
    static_assert(SPA % SPS == 0, "an A stage is a whole number of shared-memory stages");
    static_assert(DG == 1 || (DG == 3 && SPS == SPA), "grouped decode: one shared-memory stage per A block");
    constexpr int NDEC = (DG == 3) ? 12 : 8;   // decode warps 4 .. 4+NDEC-1
    constexpr int W_ACT = 4 + NDEC, W_MMA = 5 + NDEC, W_ALLOC = 6 + NDEC;
    constexpr int NPW = npw_for(DG);             // weight producers: warps W_ALLOC .. W_ALLOC+NPW-1, stage s by producer s % NPW
    constexpr int NA = (DG == 3) ? 3 : 2;      // A stages in TMEM
    constexpr int DEC_PER_BLOCK = (DG == 3) ? 4 : 8;
    constexpr int ACT_BYTES = NT * 128;
    constexpr int W_OFF = SPS * ACT_BYTES;          // stage: [activations][weights]
    constexpr uint32_t A_COL0 = 2 * NT;             // two accumulators of NT columns, then two A stages
    static_assert(A_COL0 + NA * A_COLS <= 512, "TMEM budget");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_full[MAX_STAGES], s_empty[MAX_STAGES];
    __shared__ __align__(8) uint64_t s_afull[NA], s_aempty[NA];
    __shared__ __align__(8) uint64_t s_accfull[2], s_accempty[2];
    __shared__ uint32_t s_tmem_base;
    __shared__ int s_last;
    __shared__ __align__(16) int32_t s_gs[NT];
    __shared__ __align__(16) float s_sc[NT];

    // warp index through a shuffle: provably warp-uniform for the compiler, so the role loops keep their state (counters,
    // barrier / TMEM addresses, descriptors) in uniform registers instead of converting per use (gemm_tcs.cu)
    const int warp = __shfl_sync(0xffffffffu, (int) (threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int S = p.nstages;
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;

    // let the next kernel of the stream start its own prologue / weight prefetch as early as possible
    if (p.pdl) ptx::pdl_launch_dependents();

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) {
            ptx::mbar_init(ptx::smem_u32(&s_full[s]), 2);       // the two producer warps (arrive.expect_tx each)
            ptx::mbar_init(ptx::smem_u32(&s_empty[s]), DEC_PER_BLOCK + 1);  // decode warps of the block + the MMA commit
        }
        for (int d = 0; d < NA; d++) {
            ptx::mbar_init(ptx::smem_u32(&s_afull[d]), DEC_PER_BLOCK);
            ptx::mbar_init(ptx::smem_u32(&s_aempty[d]), 1);
        }
        for (int d = 0; d < 2; d++) {
            ptx::mbar_init(ptx::smem_u32(&s_accfull[d]), 1);
            ptx::mbar_init(ptx::smem_u32(&s_accempty[d]), 4);
        }
        ptx::fence_barrier_init();
    }
    if (warp == W_ALLOC) {
        ptx::tmem_alloc(ptx::smem_u32(&s_tmem_base), 512);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = s_tmem_base;
    float *const out0 = (OUT == 0) ? p.g.out : p.g.outs[0];

    const int ntiles = p.RT * p.TT;
    const int u0 = p.whole ? (int) (((int64_t) blockIdx.x * ntiles) / gridDim.x) * p.G : (int) (((int64_t) blockIdx.x * p.U) / gridDim.x);
    const int u1 = p.whole ? (int) (((int64_t) (blockIdx.x + 1) * ntiles) / gridDim.x) * p.G
                           : (int) (((int64_t) (blockIdx.x + 1) * p.U) / gridDim.x);
    auto tile_fast = [&](int tt) -> bool { return p.allow_fast && (__ldg(p.g.act.tile_unsafe + tt) != p.g.act.epoch); };
    // optional timeline of CTA 0: (cycles since start, kind) per role, skipping the first `dbg_skip` events
    const bool dbg_on = p.dbg != nullptr && blockIdx.x == 0 && lane == 0 && (warp <= 4 || warp >= 4 + NDEC);
    const long long t_start = dbg_on ? clock64() : 0;
    int n_ev = 0, ev_role = 0;
    auto mark = [&](int kind) {
        if (dbg_on) {
            const int k = n_ev - p.dbg_skip;  // (no global load here: it would stall every marked point)
            if (k >= 0 && k < 24) {
                p.dbg[8192 + (ev_role * 24 + k) * 2] = clock64() - t_start;
                p.dbg[8192 + (ev_role * 24 + k) * 2 + 1] = kind;
            }
            ++n_ev;
        }
    };

    if (warp == W_ACT || warp >= W_ALLOC) {
        ev_role = 0;
        // ===================== TMA producers =====================
        // A warp gets one cp.async.bulk through about every memory round trip, whatever the size (~280 clk from L2, ~900 clk
        // when the source is HBM: profiles/r1_microbench_b200.txt and the small-T timelines), but warps issue
        // independently.  So the weight stream is issued by NPW warps taking the stages round robin (the ring length is a
        // multiple of NPW: every barrier has one fixed producer, no phase is skipped), activations by their own warp.
        const bool is_w = (warp >= W_ALLOC);
        const int me = warp - W_ALLOC;
        const uint64_t pol = is_w ? ptx::policy_evict_first() : ptx::policy_evict_last();
        Walk w;
        w.init(p, u0, u1);
        Ring st;
        st.init(S);
        int turn = 0;
        // weights are static: with PDL their stream starts while the predecessor (quantise) kernel is still running;
        // activations are the predecessor's output
        if (!is_w && p.pdl) ptx::pdl_wait();
        for (; w.valid(); w.next()) {
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                for (int s0 = 0; s0 < na; s0 += SPS) {
                    if (!is_w || turn == me) {
                        const int ns = min(SPS, na - s0);
                        ptx::mbar_wait(ptx::smem_u32(&s_empty[st.i]), st.ph ^ 1);
                        if (!is_w) mark(0);
                        if (ptx::elect_one()) {
                            const uint32_t full = ptx::smem_u32(&s_full[st.i]);
                            const uint32_t dst = smem_base + st.i * p.stage_bytes;
                            const int64_t kb = w.kb0 + ab0 + s0;
                            if (is_w) {
                                ptx::mbar_arrive_expect_tx(full, ns * SUB_W_BYTES);
                                ptx::bulk_g2s_hint(dst + W_OFF, p.g.w + ((int64_t) w.rt * p.KB + kb) * SUB_W_BYTES,
                                                   ns * SUB_W_BYTES, full, pol);
                            } else {
                                ptx::mbar_arrive_expect_tx(full, ns * ACT_BYTES);
                                ptx::bulk_g2s_hint(dst, p.g.act.q_sw + ((int64_t) w.tt * p.KB + kb) * ACT_BYTES, ns * ACT_BYTES,
                                                   full, pol);
                            }
                        }
                        __syncwarp();
                    }
                    st.advance();
                    if (++turn == NPW) turn = 0;
                }
            }
        }
    } else if (warp == W_MMA) {
        ev_role = 1;
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = make_idesc(NT);
        Walk w;
        w.init(p, u0, u1);
        Ring st, ar;
        st.init(S);
        ar.init(NA);
        uint32_t ndrain = 0;  // drains committed so far; accumulator buffer = ndrain % 2
        if (p.pdl) ptx::pdl_wait();  // tile_fast below is written by the predecessor (quantise) kernel
        int run_tile = -1;
        bool fast = false, open = false;
        for (; w.valid(); w.next()) {
            if (w.tile != run_tile) {
                run_tile = w.tile;
                fast = tile_fast(w.tt);
            }
            const uint32_t ab = ndrain & 1;
            if (!open) {
                ptx::mbar_wait(ptx::smem_u32(&s_accempty[ab]), ((ndrain >> 1) & 1) ^ 1);
                ptx::tc_fence_after();
            }
            const uint32_t d_tmem = tmem_base + ab * NT;
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                ptx::mbar_wait(ptx::smem_u32(&s_afull[ar.i]), ar.ph);
                mark(1);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t a_tmem = tmem_base + A_COL0 + ar.i * A_COLS;
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
                    ptx::umma_commit(ptx::smem_u32(&s_aempty[ar.i]));
                }
                __syncwarp();
                mark(8);
                for (int s0 = 0; s0 < na; s0 += SPS) st.advance();
                ar.advance();
            }
            // exact path: drain after every unit; fast path: drain after the last unit of the tile run
            if (!fast || w.u == w.run_last()) {
                if (ptx::elect_one()) ptx::umma_commit(ptx::smem_u32(&s_accfull[ab]));
                __syncwarp();
                ++ndrain;
                open = false;
            } else {
                open = true;
            }
        }
    } else if (warp >= 4 && warp < 4 + NDEC) {
        // ===================== decode: packed 2-bit codes -> u8 A operand in TMEM, 8 warps per A block ==============
        // Software-pipelined over the flattened stage sequence: the packed words of the NEXT stage are fetched (stage
        // wait, LDS, stage release) while the tcgen05.st of the current one drain, so the per-block chain is
        // A-free wait -> unpack -> st -> wait::st -> A-ready instead of also carrying the stage wait and the LDS latency.
        ev_role = 2;
        if constexpr (DG == 3) {
            // ---- grouped decode: group g = (warp - 4) / 4 owns A blocks b = g (mod 3), shared-memory stages s = g (mod 3)
            // (the ring length is a multiple of 3 and every block is exactly one stage, so a barrier is only ever waited on
            // by its own group: no phase is skipped) and A stage g.  A warp unpacks its lane quadrant of the whole block.
            const int q = (warp - 4) & 3;
            const int g = (warp - 4) >> 2;
            const int m = q * 32 + lane;
            const uint32_t taddr0 = tmem_base + ((uint32_t) (q * 32) << 16) + A_COL0 + g * A_COLS;
            Walk w;
            w.init(p, u0, u1);
            Ring st;
            st.init(S);
            int turn = 0;          // block index mod 3
            uint32_t a_ph = 0;     // phase of this group's A stage
            for (; w.valid(); w.next()) {
                for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                    if (turn == g) {
                        const int ns = min(SPA, w.nsub - ab0);
                        ptx::mbar_wait(ptx::smem_u32(&s_full[st.i]), st.ph);
                        mark(0);
                        const uint32_t w_s = smem_base + st.i * p.stage_bytes + W_OFF + m * 16;
                        uint4 cw[2 * SPA];
#pragma unroll
                        for (int c = 0; c < 2 * SPA; c++)
                            if (c < 2 * ns) cw[c] = ptx::lds128(w_s + c * TILE_BYTES);
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[st.i]));
                        ptx::mbar_wait(ptx::smem_u32(&s_aempty[g]), a_ph ^ 1);
                        mark(1);
                        ptx::tc_fence_after();
                        mark(5);
#pragma unroll
                        for (int c = 0; c < 2 * SPA; c++) {
                            if (c < 2 * ns) {
                                const uint32_t wv[4] = {cw[c].x, cw[c].y, cw[c].z, cw[c].w};
                                uint32_t r[16];
#pragma unroll
                                for (int j = 0; j < 4; j++) {
                                    r[4 * j + 0] = wv[j] & 0x03030303u;
                                    r[4 * j + 1] = __umulhi(wv[j], 1u << 30) & 0x03030303u;
                                    r[4 * j + 2] = __umulhi(wv[j], 1u << 28) & 0x03030303u;
                                    r[4 * j + 3] = __umulhi(wv[j], 1u << 26) & 0x03030303u;
                                }
                                ptx::tmem_st_32x32b_x16(taddr0 + c * 16, r);
                                if (c == 0 || c == 3) mark(6);
                            }
                        }
                        mark(2);
                        ptx::tmem_st_wait();
                        mark(4);
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_afull[g]));
                        mark(9);
                        a_ph ^= 1;
                    }
                    st.advance();
                    if (++turn == 3) turn = 0;
                }
            }
        } else {
        const int q = (warp - 4) & 3;   // TMEM lane quadrant (== warp % 4)
        const int h = (warp - 4) >> 2;  // which half of every stage this warp unpacks
        const int m = q * 32 + lane;    // weight row inside the strip == TMEM lane
        constexpr int CPS = 2 * SPS;    // 64-k chunks per stage
        constexpr int CPH = CPS / 2;    // chunks per half
        Walk w;
        w.init(p, u0, u1);
        Ring st, ar;
        st.init(S);
        ar.init(2);
        int ab0 = 0, s0 = 0;            // position of the stage whose words are in cw: unit w, A block ab0, sub-block s0
        uint4 cw[CPH];
        int ns = 0;
        auto fetch = [&]() {            // words of stage (w, ab0, s0) -> cw; the stage goes back to the producers
            ns = min(SPS, min(SPA, w.nsub - ab0) - s0);
            ptx::mbar_wait(ptx::smem_u32(&s_full[st.i]), st.ph);
            mark(0);
            const uint32_t w_s = smem_base + st.i * p.stage_bytes + W_OFF + m * 16;
#pragma unroll
            for (int c = 0; c < CPH; c++)
                if (h * CPH + c < 2 * ns) cw[c] = ptx::lds128(w_s + (h * CPH + c) * TILE_BYTES);
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[st.i]));
            st.advance();
        };
        if (w.valid()) fetch();
        while (w.valid()) {
            const int cur_s0 = s0, cur_ns = ns;
            const int na = min(SPA, w.nsub - ab0);
            const bool first = (s0 == 0), last = (s0 + SPS >= na);
            // advance the position to the next stage
            s0 += SPS;
            if (last) {
                s0 = 0;
                ab0 += SPA;
                if (ab0 >= w.nsub) { ab0 = 0; w.next(); }
            }
            if (first) {
                ptx::mbar_wait(ptx::smem_u32(&s_aempty[ar.i]), ar.ph ^ 1);
                mark(1);
                ptx::tc_fence_after();
            }
            const uint32_t taddr0 = tmem_base + ((uint32_t) (q * 32) << 16) + A_COL0 + ar.i * A_COLS;
#pragma unroll
            for (int c = 0; c < CPH; c++) {
                if (h * CPH + c < 2 * cur_ns) {
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
                    ptx::tmem_st_32x32b_x16(taddr0 + cur_s0 * 32 + (h * CPH + c) * 16, r);
                }
            }
            mark(2);
            if (w.valid()) fetch();  // next stage's words, while the stores above complete
            if (last) {
                mark(3);
                ptx::tmem_st_wait();
                mark(4);
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_afull[ar.i]));
                mark(9);
                ar.advance();
            }
        }
        }  // DG == 1
    } else if (warp < 4) {
        // ===================== epilogue =====================
        const int ew = warp;
        ev_role = 3;
        const int m = ew * 32 + lane;
        const int et = threadIdx.x;  // 0..127
        // everything the epilogue reads from global memory was written by predecessor kernels; it also owns all
        // global writes of this kernel
        if (p.pdl) ptx::pdl_wait();
        uint32_t ndrain = 0;
        // exact-path running totals: this CTA's scratch tile in the (L2-resident) workspace
        int32_t *const tots = p.g.part + ((int64_t) gridDim.x * 2 + blockIdx.x) * (NT * 128);
        Walk w;
        w.init(p, u0, u1);
        while (w.valid()) {
            // ---- one tile run: units [ra, rb] of one output tile ----
            const int ra = w.u, rb = w.run_last();
            const int g_begin = w.g, g_end = w.g + (rb - ra);
            const bool fast = tile_fast(w.tt);
            const bool complete = (g_begin == 0) && (g_end == p.G - 1);
            const int run_slot = (ra == u0) ? 0 : 1;
            const int tile = w.tile;
            const int64_t n = (int64_t) w.rt * TILE_ROWS + m;
            const int64_t t0 = (int64_t) w.tt * NT;
            const int32_t *pre = p.g.act.gpre + (t0 + et) * (p.G + 1);
            int32_t *const slot = p.g.part + ((int64_t) blockIdx.x * 2 + run_slot) * (NT * 128);
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
                                store_out<OUT>(out0, p.g.outs, p.g.n_out, (t0 + c0 + i) * p.g.ldo + n, v);
                            }
                        }
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; i++) __stcg(slot + (c0 + i) * 128 + m, vals[i]);
                }
            };

            for (int u = ra; u <= rb; u++, w.next()) {
                if (fast && u != rb) continue;  // fast path: a single drain at the end of the run
                const uint32_t ab = ndrain & 1;
                if (et < NT) s_gs[et] = __ldg(pre + w.g + 1) - __ldg(pre + (fast ? g_begin : w.g));
                ptx::bar_sync(1, 128);
                ptx::mbar_wait(ptx::smem_u32(&s_accfull[ab]), (ndrain >> 1) & 1);
                mark(0);
                ptx::tc_fence_after();
                const uint32_t ta = tmem_base + ((uint32_t) (ew * 32) << 16) + ab * NT;
                if (fast) {
                    // hot path: one drain per tile run, branch-free per value (epilogue.cuh)
                    const int64_t idx0 = t0 * p.g.ldo + n;
                    const int tvalid = (int) min((int64_t) NT, p.g.T - t0);
                    if (complete)
                        epi::drain_fast<NT, OUT, false>(ta, s_gs, s_sc, p.g.w_scale, out0, p.g.outs, p.g.n_out, idx0,
                                                        p.g.ldo, tvalid, n < p.g.N, nullptr);
                    else
                        epi::drain_fast<NT, OUT, true>(ta, s_gs, s_sc, p.g.w_scale, out0, p.g.outs, p.g.n_out, idx0,
                                                       p.g.ldo, tvalid, n < p.g.N, slot + m);
                } else {
                    // exact path: int16 wrap per K-group, running totals in the scratch tile
#pragma unroll 1
                    for (int c0 = 0; c0 < NT; c0 += 16) {
                        uint32_t v[16];
                        ptx::tmem_ld_32x32b_x16(ta + c0, v);
                        ptx::tmem_ld_wait();
                        int32_t vals[16];
#pragma unroll
                        for (int i = 0; i < 16; i++) {
                            const int32_t x = (int32_t) v[i] - s_gs[c0 + i];
                            vals[i] = (u == ra) ? wrap16(x) : tots[(c0 + i) * 128 + et] + wrap16(x);
                        }
                        if (u == rb) {
                            emit(c0, vals);
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; i++) tots[(c0 + i) * 128 + et] = vals[i];
                        }
                    }
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_accempty[ab]));
                mark(7);
                ++ndrain;
                ptx::bar_sync(1, 128);  // s_gs is rewritten by the next drain
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
                        if (n < p.g.N)
                            epi::store16<OUT>(acc, s_sc + c0, p.g.w_scale != 1.0f, p.g.w_scale, out0, p.g.outs, p.g.n_out,
                                              (t0 + c0) * p.g.ldo + n, p.g.ldo, (int) min((int64_t) 16, p.g.T - t0 - c0));
                    }
                    if (et == 0) p.g.counters[tile] = 0;
                }
            }
            ptx::bar_sync(1, 128);  // s_sc / s_last are reused by the next run
        }
    }

    // ---- teardown ----
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == W_ALLOC) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 512);
    }
}

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}

template <int NT, int SPS, int OUT, int DG>
static cudaError_t launch_out(Params &p, int grid, cudaStream_t s) {
    const int budget = env_int("VLUT_TC_SMEM_KB", 200) * 1024;
    p.stage_bytes = SPS * (NT * 128 + SUB_W_BYTES);  // multiple of 1024
    int ns = (budget - 1024) / p.stage_bytes;
    if (ns > MAX_STAGES) ns = MAX_STAGES;
    int forced = env_int("VLUT_TC_STAGES", 0);
    // the decode warps fetch the first stage of the next A block before they hand over the current one: the ring must
    // hold one stage more than an A block, or that fetch would wait for an MMA that waits for the hand-over
    constexpr int MIN_STAGES = SPA / SPS + 1;
    if (forced > 0 && forced < MIN_STAGES) forced = MIN_STAGES;
    constexpr int QUANT = (DG == 3) ? 3 : npw_for(DG);  // stage -> decode group / weight producer mapping must be static
    if (forced > 0) forced = (forced + QUANT - 1) / QUANT * QUANT;
    if (forced > 0 && forced < ns) ns = forced;
    ns -= ns % QUANT;
    if (ns < MIN_STAGES || ns < QUANT) return cudaErrorInvalidValue;
    p.nstages = ns;
    const size_t smem = (size_t) ns * p.stage_bytes + 1024;
    auto kern = kernel<NT, SPS, OUT, DG>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) grid);
    cfg.blockDim = dim3(threads_for(DG));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = p.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, p);
}

template <int NT, int SPS, int DG>
static cudaError_t launch_one(Params &p, int grid, cudaStream_t s) {
    if (p.g.n_out == 0) return launch_out<NT, SPS, 0, DG>(p, grid, s);
    return p.g.multicast ? launch_out<NT, SPS, 2, DG>(p, grid, s) : launch_out<NT, SPS, 1, DG>(p, grid, s);
}

}  // namespace tc4

bool gemm_tc4_supported(int type, int64_t K, int64_t N, int64_t T) {
    return type_valid(type) && T >= 1 && K >= 4 && N >= 1 && ceil_div(K, group_size(type)) <= 64;
}

cudaError_t launch_gemm_tc4(const GemmArgs &g, cudaStream_t s, int *launches) {
    if (g.act.sw_tile != g.act.NT) return cudaErrorInvalidValue;  // wants full-height swizzled activation tiles
    tc4::Params p{};
    p.g = g;
    p.KB = (int) (g.act.Kpad / 128);
    p.gsub = g.act.gsize / 128;
    p.G = g.act.G;
    p.RT = (int) (g.Npad / TILE_ROWS);
    const int NT = g.act.NT;
    p.TT = (int) (g.act.Tpad / NT);
    p.U = p.RT * p.TT * p.G;
    p.allow_fast = tc4::env_int("VLUT_TC_FAST", 1);
    p.pdl = tc4::env_int("VLUT_PDL", 1);
    p.dbg_skip = tc4::env_int("VLUT_TC_DEBUG_SKIP", 0);
    p.dbg = tc4::env_int("VLUT_TC_DEBUG", 0) ? debug_buffer(s, tc4::env_int("VLUT_TC_DEBUG_SKIP", 0)) : nullptr;
    // whole tiles per CTA unless stream-K's perfect balance beats the cost of fixing up split tiles (see gemm_tc5.cu)
    int grid = g.sm_count;
    const int tiles = p.RT * p.TT;
    constexpr int FIXUP_UNITS = 10;
    const int rounds = (tiles + grid - 1) / grid;
    p.whole = (rounds * p.G <= (p.U + grid - 1) / grid + FIXUP_UNITS) ? 1 : 0;
    const int forced_grid = tc4::env_int("VLUT_TC_GRID", 0);  // tests: force a grid (and with it the stream-K split paths)
    if (forced_grid > 0) p.whole = 0;
    p.whole = tc4::env_int("VLUT_TC_WHOLE", p.whole);
    if (forced_grid > 0) grid = forced_grid;
    else if (p.whole) grid = (tiles + rounds - 1) / rounds;
    if (grid > p.U) grid = p.U;
    if (grid < 1) grid = 1;
    if (launches) *launches += 1;
    const bool dg3 = tc4::env_int("VLUT_TC_DGROUPS", 1) == 3;  // grouped decode (experimental: measured within 4 % of the default)
    switch (NT) {
        case 16: return dg3 ? tc4::launch_one<16, 4, 3>(p, grid, s) : tc4::launch_one<16, 4, 1>(p, grid, s);
        case 32: return dg3 ? tc4::launch_one<32, 4, 3>(p, grid, s) : tc4::launch_one<32, 4, 1>(p, grid, s);
        case 64: return tc4::launch_one<64, 2, 1>(p, grid, s);
        case 128: return tc4::launch_one<128, 2, 1>(p, grid, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace vlut
