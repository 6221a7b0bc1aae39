#pragma once
// gemm_tc5.cuh -- VLUT_ALGO_TCP (kernel template; one translation unit per token-tile configuration instantiates it:
// gemm_tc5_nt*.cu, so that they compile in parallel; gemm_tc5.cu holds the host-side dispatch): the Vec-LUT GEMM on CTA PAIRS (cluster of 2, tcgen05 cta_group::2, M = 256).
//
// Why (profiles/r1_timeline_tc4_T512.txt): with 128 rows x 128 tokens per CTA every 512-k A block
// needs 64 KB of int8 activations + 16 KB of weights from L2 for 1024 clk of tensor-pipe time = 80 B/clk/SM, twice what
// the L2 fabric sustains per SM (~6300 B/clk chip-wide = 42.6 B/clk/SM), and the 40 KB stages leave room for only two A
// blocks of look-ahead: the single-CTA kernels are shared-memory-fill bound from T = 256 up.  A CTA pair halves the
// activation traffic: one tcgen05.mma.cta_group::2 multiplies 256 weight rows (128 per CTA, unpacked into each CTA's own
// TMEM) with a 128-token activation tile of which each CTA loads and holds only HALF (64 tokens); the tensor cores read
// the peer's half over the pair's shared-memory link.  Per CTA and A block: 32 KB + 16 KB = 48 KB (47 B/clk), stages of
// 24 KB, eight stages = four A blocks of look-ahead.
//
// Roles per CTA (same warp layout as gemm_tc4.cu):
//   warps 0-3 epilogue, 4-11 decode, 12 activation producer, 13 MMA issuer (even CTA only), 14 TMEM allocator, then
//   weight producer.
// Cross-CTA protocol: every CTA has its own stage ring (full/empty), A ring (aempty) and accumulator barriers (accfull);
// the even CTA ("leader") owns afull[] (16 arrivals: 8 decode warps of each CTA -- a decode warp has waited for its CTA's
// stage, so A-ready implies that CTA's half of the activations has landed) and accempty[] (8 arrivals: 4 epilogue warps of
// each CTA).  The leader's commits are multicast to both CTAs.
#include <type_traits>
#include "epilogue.cuh"
#include "ptx.cuh"
#include "vlut_common.cuh"

namespace vlut {

namespace tc5 {

constexpr int THREADS = 480;          // warps 0-3 epilogue, 4-11 decode, 12 activation producer, 13 MMA issuer, 14 TMEM + weight producer
#ifndef VLUT_TC_EPI2
#define VLUT_TC_EPI2 0
#endif
// VLUT_TC_EPI2=1 (compile time): a SECOND epilogue warpgroup (warps 15-18) for the direct-store flavour, the two groups
// splitting every tile's columns.  Motivation: the drain of a tile takes a warp ~8200 clk inside the main loop (as long as
// the tile's MMAs) and ~3000 clk alone.  Measured (bit-exact, 631 tests): 608 threads cap the kernel at 96 registers and
// add four more warps to the schedulers the decode warps run on -- T = 512 25.6 -> 26.3 us, the 256-token tile at
// T = 1024 49.2 -> 63.4 us.  Off.
__host__ __device__ constexpr int epi_groups(int out) { return (VLUT_TC_EPI2 && out == 0) ? 2 : 1; }
__host__ __device__ constexpr int threads_for(int out) { return THREADS + (epi_groups(out) - 1) * 128; }
constexpr int MAX_NA = 4;          // A stages in TMEM (ring)
constexpr int MAX_STAGES = 12;
constexpr int SUB_W_BYTES = 4096;

struct Params {
    GemmArgs g;
    int KB, G, gsub, RT, RP, TT, U;
    int nstages, stage_bytes;
    int allow_fast, pdl;
    int split_last;  // experiment: walk the whole tiles before the split units
    int coop_fix;    // hybrid: the contributors of a split tile share its fix-up (VLUT_TC_COOP, default 1)
    int coop_timeout;   // ... and wait at most this many clk for each other (VLUT_TC_COOP_TIMEOUT)
    int uniform;  // every K-group unit has gsub sub-blocks (KB % gsub == 0): the decode warps run their lean loop
    // FLAT walk (shapes whose K-groups do not fit the lean loops: 640-k groups, K not a multiple of the group size).  On the
    // fast path (no int16 wrap possible) K-group boundaries mean nothing: the whole K extent accumulates anyway.  When the
    // quantise kernel reports that NO token tile of this call is unsafe, every role walks K in uniform A blocks instead:
    // unit = one A block of SPA sub-blocks (the last one of a tile may be shorter), Gf units per tile.
    int flat_ok, Gf, Uf, wholef;
    int uniform_knob;   // VLUT_TC_LEAN
    int whole;  // 1: whole output tiles only (no fix-up); 0: stream-K over (tile, K-group) units; 2: hybrid (whole tiles + split remainder)
    int abl;  // timing ablations (VLUT_TC_ABL, results become wrong): 1 no tcgen05.st, 2 no MMA, 4 no wait::st, 8 no output stores (exact path), 16 no fast-path drain, 32 no activation copies, 64 no weight copies, 256 no LDS / unpack, 512 no stage ring at all (producers idle, decode does not wait for stages)
    long long *dbg;
    int dbg_skip, dbg_cta;
    alignas(64) CUtensorMap out_map;   // OUT == 4: the output [T][ldo] f32 as a 2-D tensor, box 128 rows x 16 tokens (TMA stores)
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
// kind::i8, A = u8 (K-major, TMEM), B = s8 (K-major), D = s32, M = 256 across the pair, N = n
__host__ __device__ constexpr uint32_t make_idesc(int n) {
    return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (256 >> 4) << 24);
}

// units of this PAIR: unit = (row-pair tile rp, token tile tt, K-group g), ordered [rp][tt][g].  A pair walks up to two
// unit ranges: first its share [b0, b1) of the SPLIT units (tiles shared with other pairs, fixed up through the workspace),
// then its WHOLE tiles [a0, a1).  Split units go first so that the last arriver's fix-up runs under the following tiles' MMAs.
struct Ranges {
    int a0, a1, b0, b1;
};
struct Walk {
    int u, u1, n0, n1, G, TT, gsub, KB;
    int tile, g, rp, tt, kb0, nsub;
    __device__ __forceinline__ void init(const Params &p, const Ranges &r, int G_, int gsub_) {
        G = G_; TT = p.TT; gsub = gsub_; KB = p.KB;
        int f0 = r.b0, f1 = r.b1, s0 = r.a0, s1 = r.a1;
        if (p.split_last) { f0 = r.a0; f1 = r.a1; s0 = r.b0; s1 = r.b1; }
        if (f0 < f1) { u = f0; u1 = f1; n0 = s0; n1 = s1; }
        else         { u = s0; u1 = s1; n0 = n1 = 0; }
        locate();
    }
    __device__ __forceinline__ void locate() {
        tile = u / G; g = u - tile * G;
        rp = tile / TT; tt = tile - rp * TT;
        refresh();
    }
    __device__ __forceinline__ void refresh() { kb0 = g * gsub; nsub = min(gsub, KB - kb0); }
    __device__ __forceinline__ bool valid() const { return u < u1; }
    __device__ __forceinline__ int run_last() const { return min(u1 - 1, u - g + G - 1); }
    __device__ __forceinline__ void next() {
        ++u;
        if (u == u1 && n0 < n1) {  // on to the second range
            u = n0; u1 = n1; n0 = n1;
            locate();
            return;
        }
        if (++g == G) { g = 0; ++tile; if (++tt == TT) { tt = 0; ++rp; } }
        refresh();
    }
};

struct Ring {
    int i, n;
    uint32_t ph;
    __device__ __forceinline__ void init(int n_) { i = 0; n = n_; ph = 0; }
    __device__ __forceinline__ void advance() { if (++i == n) { i = 0; ph ^= 1; } }
};

// NT: tokens per tile (N of the pair MMA); every CTA stages NT/2 of them.  OUT: epilogue store flavour.
// SPA: 128-k sub-blocks per A stage (= per decode -> MMA hand-off); NA: A stages in TMEM.  (SPA, NA) = (4, 2): two hand-offs of
// 512 k in flight; (2, 4): four of 256 k -- the in-kernel timeline (profiles/r2_timeline_tcp.txt) showed the issuer waiting
// ~700 clk per 512-k block for the decode -> st -> wait::st -> remote-arrive chain with only two stages to hide it behind.
template <int NT, int SPS, int OUT, int SPA, int NA>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(threads_for(OUT), 1) kernel(const __grid_constant__ Params p) {
    constexpr int EG = epi_groups(OUT);              // epilogue warpgroups
    constexpr int ET = 128 * EG;                     // epilogue threads (named barrier 1)
    constexpr int A_COLS = SPA * 32;                 // TMEM columns of one A stage
    static_assert(NA <= MAX_NA, "A ring");
    static_assert(SPA % SPS == 0, "an A stage is a whole number of shared-memory stages");
    static_assert(NT % 32 == 0 && NT >= 32 && NT <= 256, "pair MMA with A in TMEM: N multiple of 32");
    constexpr int DW = 8;                            // decode warps per A block
    constexpr int NH = NT / 2;                       // tokens staged by this CTA
    constexpr int ACT_BYTES = NH * 128;              // one 128-k sub-block of this CTA's half
    constexpr int W_OFF = SPS * ACT_BYTES;           // stage: [activations][weights]
    // TMEM: NACC accumulators of NT columns, then two A stages of 128 columns.  NT = 256 (one decoded A stage multiplies TWO
    // 128-token halves: the weights are unpacked T/256 instead of T/128 times and every decode -> MMA hand-off feeds 2048
    // clk of tensor-pipe time) leaves room for ONE accumulator: the drain of a tile is not overlapped with the next tile's
    // MMAs, but its loads and unpacks are (both A stages and the whole shared-memory ring refill meanwhile).
    constexpr int NACC = (2 * NT + NA * A_COLS <= 512) ? 2 : 1;
    constexpr uint32_t A_COL0 = NACC * NT;
    static_assert(A_COL0 + NA * A_COLS <= 512, "TMEM budget");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_full[MAX_STAGES], s_empty[MAX_STAGES];
    __shared__ __align__(8) uint64_t s_afull[MAX_NA], s_aempty[MAX_NA];
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
    const uint32_t rank = ptx::cluster_ctarank();    // 0 = leader (issues the MMAs), 1 = peer
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (p.pdl) ptx::pdl_launch_dependents();

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) {
            ptx::mbar_init(ptx::smem_u32(&s_full[s]), 2);       // the two producer warps (arrive.expect_tx each)
            ptx::mbar_init(ptx::smem_u32(&s_empty[s]), DW + 1);  // the 4 local decode warps of the block's group + the leader's multicast commit
        }
        for (int d = 0; d < NA; d++) {
            ptx::mbar_init(ptx::smem_u32(&s_afull[d]), 2 * DW);     // used in the leader: the block's decode group in both CTAs
            ptx::mbar_init(ptx::smem_u32(&s_aempty[d]), 1);
        }
        for (int d = 0; d < 2; d++) {
            ptx::mbar_init(ptx::smem_u32(&s_accfull[d]), 1);
            ptx::mbar_init(ptx::smem_u32(&s_accempty[d]), 8 * EG);   // used in the leader: epilogue warps of both CTAs
        }
        ptx::fence_barrier_init();
    }
    if (warp == 14) {
        ptx::tmem_alloc_pair(ptx::smem_u32(&s_tmem_base), 512);
        ptx::tmem_relinquish_pair();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();  // the peer's barriers are initialised before anybody arrives on them
    ptx::tc_fence_after();
    const uint32_t tmem_base = s_tmem_base;
    float *const out0 = (OUT == 0 || OUT == 4) ? p.g.out : p.g.outs[0];
    static_assert(OUT >= 0 && OUT <= 4, "epilogue store flavour");
    constexpr bool STAGED = (OUT == 3 || OUT == 4);   // tiles leave through shared memory + bulk async copies
    constexpr int SO = (OUT == 4) ? 0 : OUT;          // flavour of the direct stores (exact path)

    // walk shape: the K-group units of the op, or (flat) one A block per unit when no tile of this call can wrap
    int shG = p.G, shgsub = p.gsub, shU = p.U, shwhole = p.whole;
    bool flat = false;
    if (p.flat_ok) {
        if (p.pdl) ptx::pdl_wait();   // every role, the weight producer included: the flag is the quantise kernel's output
        flat = p.allow_fast && (__ldg(p.g.act.tile_unsafe + (p.g.act.Tpad >> 4)) != p.g.act.epoch);
        if (flat) { shG = p.Gf; shgsub = SPA; shU = p.Uf; shwhole = p.wholef; }
    }
    const int ntiles = p.RP * p.TT;
    // p.whole: 1 = whole tiles only (uneven tile counts per pair), 0 = stream-K (every unit is a split unit),
    // 2 = hybrid: floor(ntiles / npairs) whole tiles per pair, the units of the remaining tiles split evenly over all pairs
    const int full = (shwhole == 2) ? ntiles / npairs : 0;
    const int sbase = npairs * full * shG;   // first split unit
    const int SU = shU - sbase;              // split units
    Ranges rg;
    if (shwhole == 1) {
        rg.a0 = (int) (((int64_t) pair * ntiles) / npairs) * shG;
        rg.a1 = (int) (((int64_t) (pair + 1) * ntiles) / npairs) * shG;
        rg.b0 = rg.b1 = 0;
    } else {
        rg.a0 = pair * full * shG;
        rg.a1 = (pair + 1) * full * shG;
        rg.b0 = sbase + (int) (((int64_t) pair * SU) / npairs);
        rg.b1 = sbase + (int) (((int64_t) (pair + 1) * SU) / npairs);
    }
    // lean role loops (issuer, decode): every K-group unit has p.gsub sub-blocks and every A block is full, and no ablation
    // asks for the generic loops
    // (flat: one block per unit, the last unit of a tile possibly short -- handled by the lean loops when a block is one
    // shared-memory stage)
    const bool lean = (flat ? (SPS == SPA && p.uniform_knob) : (p.uniform && (p.gsub % SPA == 0))) && !(p.abl & (256 | 512));
    const int tail_na = p.KB - (shG - 1) * shgsub - (shgsub / SPA - 1) * SPA;   // sub-blocks of a tile's LAST A block (lean shapes)
    auto tile_fast = [&](int tt) -> bool { return p.allow_fast && (__ldg(p.g.act.tile_unsafe + tt) != p.g.act.epoch); };
    // optional timeline of CTA 0: (cycles since start, kind) per role, skipping the first `dbg_skip` events
    const bool dbg_on = p.dbg != nullptr && (int) blockIdx.x == p.dbg_cta && lane == 0;
    const long long t_start = dbg_on ? clock64() : 0;
    const long long t_start0 = p.dbg != nullptr ? clock64() : 0;
    long long t_ns0 = 0;   // wall clock (globaltimer, ns) next to the SM clock: their ratio is the SM frequency under load
    if (p.dbg != nullptr) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_ns0));
    int n_ev = 0, ev_role = 0;
    auto mark = [&](int kind) {
        if (dbg_on && ev_role < 4) {
            const int k = n_ev - p.dbg_skip;  // (no global load here: it would stall every marked point)
            if (k >= 0 && k < 24) {
                p.dbg[8192 + (ev_role * 24 + k) * 2] = clock64() - t_start;
                p.dbg[8192 + (ev_role * 24 + k) * 2 + 1] = kind;
            }
            ++n_ev;
        }
    };

    // per-role wait accounting (VLUT_TC_DEBUG): cycles spent inside each kind of barrier wait, summed in registers and
    // written once at role exit -- no marks, so the timed run keeps the production instruction stream except for two
    // clock reads per wait.  Layout: dbg[(cta * 4 + role) * 8 + k], k = 7: life of the role.
    const bool cnt_on = p.dbg != nullptr;
    long long tw[4] = {0, 0, 0, 0};
    auto twait = [&](uint32_t bar, uint32_t parity, int k) {
        if (cnt_on) {
            const long long t = clock64();
            ptx::mbar_wait(bar, parity);
            tw[k] += clock64() - t;
        } else {
            ptx::mbar_wait(bar, parity);
        }
    };
    auto tflush = [&](int role, int k0) {
        if (cnt_on && blockIdx.x < 256) {
            long long *o = p.dbg + ((int) blockIdx.x * 4 + role) * 8;
            o[k0] = tw[0]; o[k0 + 1] = tw[1]; o[k0 + 2] = tw[2];
            if (k0 == 0) {
                long long t_ns1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_ns1));
                o[7] = clock64() - t_start0; o[6] = tw[3]; o[5] = t_ns1 - t_ns0;
            }
        }
    };
    if (warp == 12 || warp == 14) {
        ev_role = 0;
        // ===================== two TMA producers: warp 14 streams this CTA's weight rows, warp 12 its half of the
        // activation tile.  One cp.async.bulk costs its issuing thread ~280 clk whatever its size
        // (profiles/r1_microbench_b200.txt): few, large copies, issued from two warps. ==============
        const bool is_w = (warp == 14);
        const uint64_t pol = is_w ? ptx::policy_evict_first() : ptx::policy_evict_last();
        Walk w;
        w.init(p, rg, shG, shgsub);
        Ring st;
        st.init(S);
        // weights are static: with PDL their stream starts while the predecessor (quantise) kernel is still running;
        // activations are the predecessor's output
        if (!is_w && p.pdl) ptx::pdl_wait();
        for (; w.valid() && !(p.abl & 512); w.next()) {
            const int rt = min(2 * w.rp + (int) rank, p.RT - 1);  // odd strip count: the last peer re-reads a valid strip
            for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                const int na = min(SPA, w.nsub - ab0);
                for (int s0 = 0; s0 < na; s0 += SPS) {
                    const int ns = min(SPS, na - s0);
                    twait(ptx::smem_u32(&s_empty[st.i]), st.ph ^ 1, 0);
                    if (!is_w) mark(0);
                    if (ptx::elect_one()) {
                        const uint32_t full = ptx::smem_u32(&s_full[st.i]);
                        const uint32_t dst = smem_base + st.i * p.stage_bytes;
                        const int64_t kb = w.kb0 + ab0 + s0;
                        if (p.abl & (is_w ? 64 : 32)) {
                            ptx::mbar_arrive(full);   // timing ablation: this operand stream is not fetched
                        } else if (is_w) {
                            ptx::mbar_arrive_expect_tx(full, ns * SUB_W_BYTES);
                            ptx::bulk_g2s_hint(dst + W_OFF, p.g.w + ((int64_t) rt * p.KB + kb) * SUB_W_BYTES, ns * SUB_W_BYTES,
                                               full, pol);
                        } else {
                            // q_sw holds swizzled tiles of NH tokens: this CTA's half is contiguous over kb
                            ptx::mbar_arrive_expect_tx(full, ns * ACT_BYTES);
                            ptx::bulk_g2s_hint(dst, p.g.act.q_sw + (((int64_t) (2 * w.tt + rank) * p.KB) + kb) * ACT_BYTES,
                                               ns * ACT_BYTES, full, pol);
                        }
                    }
                    __syncwarp();
                    st.advance();
                }
            }
        }
        if (!is_w && lane == 0) tflush(0, 0);
    } else if (warp == 13) {
        ev_role = 1;
        // ===================== MMA issuer (leader CTA only): one M=256 MMA feeds both CTAs' accumulators ============
        // ONE thread runs the whole role (elected once, not per block): the issuer is a serial instruction stream, and the
        // in-kernel timeline (profiles/r2_timeline_tcp_issuer.txt) showed ~450-640 clk of it per A block plus ~320 clk per
        // K-group unit (per-block election + __syncwarp, ring copies, descriptors rebuilt per sub-block, loop bounds) against
        // 512 clk of tensor-pipe time per 256-k block.  Here a full block is one unrolled burst: operands are the stage's
        // descriptor word plus compile-time offsets, barrier addresses are ring index * 8 off a base kept in registers.
        if (rank == 0 && ptx::elect_one()) {
            constexpr uint32_t idesc = make_idesc(NT);
            constexpr uint32_t DESC_HI = (uint32_t) ((1024 >> 4) | (1u << 14) | (2u << 29));   // SBO 1024 B, version 1, SWIZZLE_128B
            constexpr uint32_t DESC_LBO = 1u << 16;
            Walk w;
            w.init(p, rg, shG, shgsub);
            int st_i = 0, ar_i = 0;
            uint32_t ar_ph = 0;
            const uint32_t empty0 = ptx::smem_u32(&s_empty[0]), afull0 = ptx::smem_u32(&s_afull[0]);
            const uint32_t aempty0 = ptx::smem_u32(&s_aempty[0]);
            const uint32_t stage_lo0 = (smem_base & 0x3FFFFu) >> 4, stage_step = (uint32_t) p.stage_bytes >> 4;
            uint32_t stage_lo = stage_lo0;   // descriptor start-address field of stage st_i
            uint32_t ndrain = 0;  // drains committed so far; accumulator buffer = ndrain % NACC
            if (p.pdl) ptx::pdl_wait();  // tile_fast below is written by the predecessor (quantise) kernel
            int run_tile = -1;
            bool fast = false, open = false;
            const bool no_mma = (p.abl & 2) != 0;
            // (one body, compiled twice: the flat walk's short-block bookkeeping must not cost the K-group walk anything --
            //  a few instructions per block in this loop are ~2 % of the kernel)
            auto lean_issue = [&](auto flat_c) {
                constexpr bool FLAT = decltype(flat_c)::value;
                // ---- lean loop: every A block is full; the unit walk happens once per TILE RUN, not per block ----
                const uint32_t nbpu = (uint32_t) (shgsub / SPA);   // A blocks per unit
                while (w.valid()) {
                    const int nu = w.run_last() - w.u + 1;         // units of this tile run
                    const bool fast_run = tile_fast(w.tt);
                    // fast path: one accumulation segment (a single drain) per run; exact path: one per unit
                    const int nseg = fast_run ? 1 : nu;
                    const uint32_t seg_blocks = (fast_run ? (uint32_t) nu : 1u) * nbpu;
                    // blocks of this run until (and including) the tile's last block, the only one that can be short
                    [[maybe_unused]] uint32_t to_tail = (uint32_t) (shG - w.g) * nbpu;
                    for (int seg = 0; seg < nseg; seg++) {
                        const uint32_t ab = ndrain % NACC;
                        ptx::mbar_wait(ptx::smem_u32(&s_accempty[ab]), ((ndrain / NACC) & 1) ^ 1);
                        ptx::tc_fence_after();
                        const uint32_t d_tmem = tmem_base + ab * NT;
                        uint32_t acc = 0u;
                        for (uint32_t b = 0; b < seg_blocks; b++) {
                            ptx::mbar_wait(afull0 + 8u * (uint32_t) ar_i, ar_ph);
                            ptx::tc_fence_after();
                            const uint32_t a_tmem = tmem_base + A_COL0 + (uint32_t) ar_i * A_COLS;
                            bool short_blk = false;
                            if constexpr (FLAT) short_blk = (--to_tail == 0 && tail_na != SPA);
                            if (FLAT && short_blk) {
                                // the tile's short last block (flat walk; SPS == SPA: one stage): tail_na sub-blocks
                                if (!no_mma) {
                                    for (int sb = 0; sb < tail_na; sb++) {
#pragma unroll
                                        for (int ks = 0; ks < 4; ks++) {
                                            const uint64_t b_desc = ((uint64_t) DESC_HI << 32) |
                                                                    (uint64_t) (DESC_LBO | (stage_lo + (uint32_t) (sb * (ACT_BYTES >> 4) + ks * 2)));
                                            ptx::umma_i8_ts_pair(d_tmem, a_tmem + sb * 32 + ks * 8, b_desc, idesc,
                                                                 (sb == 0 && ks == 0) ? acc : 1u);
                                        }
                                    }
                                }
                                ptx::umma_commit_pair(aempty0 + 8u * (uint32_t) ar_i, 3);
                                ptx::umma_commit_pair(empty0 + 8u * (uint32_t) st_i, 3);
                                if (++st_i == S) { st_i = 0; stage_lo = stage_lo0; } else { stage_lo += stage_step; }
                            } else {
#pragma unroll
                            for (int s0 = 0; s0 < SPA; s0 += SPS) {
                                if (!no_mma) {
#pragma unroll
                                    for (int sb = 0; sb < SPS; sb++) {
#pragma unroll
                                        for (int ks = 0; ks < 4; ks++) {
                                            const uint64_t b_desc = ((uint64_t) DESC_HI << 32) |
                                                                    (uint64_t) (DESC_LBO | (stage_lo + (uint32_t) (sb * (ACT_BYTES >> 4) + ks * 2)));
                                            ptx::umma_i8_ts_pair(d_tmem, a_tmem + (s0 + sb) * 32 + ks * 8, b_desc, idesc,
                                                                 (s0 == 0 && sb == 0 && ks == 0) ? acc : 1u);
                                        }
                                    }
                                }
                                // the A stage first: the decode warps' wait for it is on the hand-off loop's critical path
                                if (s0 + SPS >= SPA) ptx::umma_commit_pair(aempty0 + 8u * (uint32_t) ar_i, 3);
                                ptx::umma_commit_pair(empty0 + 8u * (uint32_t) st_i, 3);
                                if (++st_i == S) { st_i = 0; stage_lo = stage_lo0; } else { stage_lo += stage_step; }
                            }
                            }
                            acc = 1u;
                            if (++ar_i == NA) { ar_i = 0; ar_ph ^= 1u; }
                        }
                        ptx::umma_commit_pair(ptx::smem_u32(&s_accfull[ab]), 3);
                        ++ndrain;
                    }
                    for (int i = 0; i < nu; i++) w.next();
                }
            };
            if (lean) {
                if (flat) lean_issue(std::true_type{}); else lean_issue(std::false_type{});
            } else
            for (; w.valid(); w.next()) {
                if (w.tile != run_tile) {
                    run_tile = w.tile;
                    fast = tile_fast(w.tt);
                }
                const uint32_t ab = ndrain % NACC;
                if (!open) {
                    twait(ptx::smem_u32(&s_accempty[ab]), ((ndrain / NACC) & 1) ^ 1, 2);
                    ptx::tc_fence_after();
                }
                const uint32_t d_tmem = tmem_base + ab * NT;
                uint32_t acc_first = open ? 1u : 0u;   // accumulate flag of the first MMA of this unit
                for (int ab0 = 0; ab0 < w.nsub; ab0 += SPA) {
                    const int na = min(SPA, w.nsub - ab0);
                    twait(afull0 + 8u * (uint32_t) ar_i, ar_ph, 1);
                    mark(1);
                    ptx::tc_fence_after();
                    const uint32_t a_tmem = tmem_base + A_COL0 + (uint32_t) ar_i * A_COLS;
                    if (na == SPA) {
                        // full block: SPA / SPS stages, every MMA operand a compile-time offset
#pragma unroll
                        for (int s0 = 0; s0 < SPA; s0 += SPS) {
#pragma unroll
                            for (int sb = 0; sb < SPS; sb++) {
#pragma unroll
                                for (int ks = 0; ks < 4; ks++) {
                                    const uint64_t b_desc = ((uint64_t) DESC_HI << 32) |
                                                            (uint64_t) (DESC_LBO | (stage_lo + (uint32_t) (sb * (ACT_BYTES >> 4) + ks * 2)));
                                    if (!no_mma)
                                        ptx::umma_i8_ts_pair(d_tmem, a_tmem + (s0 + sb) * 32 + ks * 8, b_desc, idesc,
                                                             (s0 == 0 && sb == 0 && ks == 0) ? acc_first : 1u);
                                }
                            }
                            // the A stage first: the decode warps' wait for it is on the hand-off loop's critical path
                            if (s0 + SPS >= SPA) ptx::umma_commit_pair(aempty0 + 8u * (uint32_t) ar_i, 3);
                            ptx::umma_commit_pair(empty0 + 8u * (uint32_t) st_i, 3);
                            if (++st_i == S) { st_i = 0; stage_lo = stage_lo0; } else { stage_lo += stage_step; }
                        }
                    } else {
                        for (int s0 = 0; s0 < na; s0 += SPS) {
                            const int ns = min(SPS, na - s0);
                            for (int sb = 0; sb < ns; sb++) {
#pragma unroll
                                for (int ks = 0; ks < 4; ks++) {
                                    const uint64_t b_desc = ((uint64_t) DESC_HI << 32) |
                                                            (uint64_t) (DESC_LBO | (stage_lo + (uint32_t) (sb * (ACT_BYTES >> 4) + ks * 2)));
                                    if (!no_mma)
                                        ptx::umma_i8_ts_pair(d_tmem, a_tmem + (s0 + sb) * 32 + ks * 8, b_desc, idesc,
                                                             (s0 == 0 && sb == 0 && ks == 0) ? acc_first : 1u);
                                }
                            }
                            ptx::umma_commit_pair(empty0 + 8u * (uint32_t) st_i, 3);
                            if (++st_i == S) { st_i = 0; stage_lo = stage_lo0; } else { stage_lo += stage_step; }
                        }
                        ptx::umma_commit_pair(aempty0 + 8u * (uint32_t) ar_i, 3);
                    }
                    mark(8);
                    acc_first = 1u;
                    if (++ar_i == NA) { ar_i = 0; ar_ph ^= 1u; }
                }
                // exact path: drain after every unit; fast path: drain after the last unit of the tile run
                if (!fast || w.u == w.run_last()) {
                    ptx::umma_commit_pair(ptx::smem_u32(&s_accfull[ab]), 3);
                    ++ndrain;
                    open = false;
                } else {
                    open = true;
                }
            }
            tflush(1, 0);
        }
        __syncwarp();
    } else if (warp >= 4 && warp < 12) {
        // ===================== decode: packed 2-bit codes -> u8 A operand in TMEM, 8 warps per A block ==============
        // The hand-off loop (A stage free -> decode -> A ready -> MMAs issued -> commit -> A stage free) is a chain of serial
        // instruction streams, and with two A stages its latency IS the kernel's period (scripts/abl_tcp.py: the barrier
        // skeleton alone ran at ~1200 clk per 512-k block against 1024 clk of tensor-pipe time; every instruction added to
        // a role loop cost ~5 clk of period).  So everything that does not need the A stage happens BEFORE the wait for it:
        // the packed words of a stage are fetched (stage wait, LDS, stage release) and unpacked into registers while the
        // previous block is still being multiplied; after the A-free wait only the tcgen05.st, wait::st and the remote
        // arrive remain (~250 clk instead of ~850).
        ev_role = 2;
        const int q = (warp - 4) & 3;   // TMEM lane quadrant (== warp % 4)
        const int h = (warp - 4) >> 2;  // which half of every stage this warp unpacks
        const int m = q * 32 + lane;    // weight row inside the strip == TMEM lane
        constexpr int CPS = 2 * SPS;    // 64-k chunks per stage
        constexpr int CPH = CPS / 2;    // chunks per half
        const uint32_t afull_leader0 = ptx::mapa(ptx::smem_u32(&s_afull[0]), 0);   // the ring's barriers are 8 bytes apart
        const uint32_t stage_bytes = (uint32_t) p.stage_bytes;
        const uint32_t w_s0 = smem_base + W_OFF + m * 16 + h * CPH * TILE_BYTES;
        const uint32_t taddr_base = tmem_base + ((uint32_t) (q * 32) << 16) + A_COL0 + h * CPH * 16;
        const bool do_st = !(p.abl & 1), do_wait = !(p.abl & 4);
        Walk w;
        w.init(p, rg, shG, shgsub);
        Ring st, ar;
        st.init(S);
        ar.init(NA);
        int ab0 = 0, s0 = 0;            // position of the stage whose unpacked words are in r: unit w, A block ab0, sub-block s0
        uint32_t r[CPH][16];
        int ns = 0;
        auto fetch_unpack = [&]() {     // stage (w, ab0, s0) -> r (unpacked); the stage goes back to the producers
            ns = min(SPS, min(SPA, w.nsub - ab0) - s0);
            if (!(p.abl & 512)) twait(ptx::smem_u32(&s_full[st.i]), st.ph, 0);
            const uint32_t w_s = w_s0 + st.i * stage_bytes;
            uint4 cw[CPH];
            if (p.abl & 256) { st.advance(); if (!(p.abl & 512) && lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[(st.i + S - 1) % S])); return; }
#pragma unroll
            for (int c = 0; c < CPH; c++)
                if (h * CPH + c < 2 * ns) cw[c] = ptx::lds128(w_s + c * TILE_BYTES);
            __syncwarp();
            if (!(p.abl & 512) && lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_empty[st.i]));
            st.advance();
#pragma unroll
            for (int c = 0; c < CPH; c++) {
                const uint32_t wv[4] = {cw[c].x, cw[c].y, cw[c].z, cw[c].w};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    // shifts as multiply-high on the FMA pipe, masks on the ALU pipe: both pipes work
                    r[c][4 * j + 0] = wv[j] & 0x03030303u;
                    r[c][4 * j + 1] = __umulhi(wv[j], 1u << 30) & 0x03030303u;
                    r[c][4 * j + 2] = __umulhi(wv[j], 1u << 28) & 0x03030303u;
                    r[c][4 * j + 3] = __umulhi(wv[j], 1u << 26) & 0x03030303u;
                }
                epi::reg_fence(r[c]);   // unpacked HERE, not sunk below the A-free wait
            }
        };
        auto lean_decode = [&](auto flat_c) {
            constexpr bool FLAT = decltype(flat_c)::value;
            // ---- lean loop: decode needs neither tiles nor units, only the number of (full) A blocks of this pair ----
            constexpr int SB = SPA / SPS;                        // shared-memory stages per A block
            const uint32_t nbpu = (uint32_t) (shgsub / SPA);     // A blocks per unit
            const uint32_t NB = (uint32_t) ((rg.a1 - rg.a0) + (rg.b1 - rg.b0)) * nbpu;
            // which blocks are a tile's short last block (flat walk only, SB == 1): blocks until the next tile end, per range
            const bool has_tail = (tail_na != SPA);
            const int r0b = p.split_last ? rg.a0 : rg.b0, r0e = p.split_last ? rg.a1 : rg.b1;   // first range walked
            const int r1b = p.split_last ? rg.b0 : rg.a0;                                        // second range
            const uint32_t nb_first = (uint32_t) (r0e - r0b) * nbpu;
            auto blocks_to_tail = [&](int u) { return (uint32_t) (shG - (u % shG)) * nbpu; };
            const uint32_t full0 = ptx::smem_u32(&s_full[0]), empty0 = ptx::smem_u32(&s_empty[0]);
            const uint32_t aempty0 = ptx::smem_u32(&s_aempty[0]);
            uint32_t st_i = 0, st_ph = 0, ar_i = 0, ar_ph = 0;
            // (a short block's missing chunks are loaded and unpacked as whatever the stage holds, and never stored)
            auto fetch_unpack_lean = [&]() {
                ptx::mbar_wait(full0 + 8u * st_i, st_ph);
                const uint32_t w_s = w_s0 + st_i * stage_bytes;
                uint4 cw[CPH];
#pragma unroll
                for (int c = 0; c < CPH; c++) cw[c] = ptx::lds128(w_s + c * TILE_BYTES);
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(empty0 + 8u * st_i);
                if (++st_i == (uint32_t) S) { st_i = 0; st_ph ^= 1u; }
#pragma unroll
                for (int c = 0; c < CPH; c++) {
                    const uint32_t wv[4] = {cw[c].x, cw[c].y, cw[c].z, cw[c].w};
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        r[c][4 * j + 0] = wv[j] & 0x03030303u;
                        r[c][4 * j + 1] = __umulhi(wv[j], 1u << 30) & 0x03030303u;
                        r[c][4 * j + 2] = __umulhi(wv[j], 1u << 28) & 0x03030303u;
                        r[c][4 * j + 3] = __umulhi(wv[j], 1u << 26) & 0x03030303u;
                    }
                    epi::reg_fence(r[c]);
                }
            };
            uint32_t more = NB * SB;                             // stages not fetched yet
            if (more) { --more; fetch_unpack_lean(); }
            [[maybe_unused]] uint32_t to_tail = NB ? blocks_to_tail(nb_first ? r0b : r1b) : 0;
            for (uint32_t blk = 0; blk < NB; blk++) {
                bool short_blk = false;
                if constexpr (FLAT) {
                    if (blk == nb_first && nb_first) to_tail = blocks_to_tail(r1b);   // on to the second range
                    short_blk = (--to_tail == 0) && has_tail;
                    if (to_tail == 0) to_tail = (uint32_t) shG * nbpu;
                }
                ptx::mbar_wait(aempty0 + 8u * ar_i, ar_ph ^ 1u);
                ptx::tc_fence_after();
                const uint32_t taddr0 = taddr_base + ar_i * A_COLS;
#pragma unroll
                for (int sg = 0; sg < SB; sg++) {
                    if (do_st) {
#pragma unroll
                        for (int c = 0; c < CPH; c++)
                            if (!FLAT || !short_blk || h * CPH + c < 2 * tail_na) ptx::tmem_st_32x32b_x16(taddr0 + sg * SPS * 32 + c * 16, r[c]);
                    }
                    if (sg == SB - 1) {
                        if (do_wait) ptx::tmem_st_wait();
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive_cluster(afull_leader0 + 8u * ar_i);
                        if (++ar_i == NA) { ar_i = 0; ar_ph ^= 1u; }
                    }
                    if (more) { --more; fetch_unpack_lean(); }
                }
            }
        };
        if (lean) {
            if (flat) lean_decode(std::true_type{}); else lean_decode(std::false_type{});
        } else {
        if (w.valid()) fetch_unpack();
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
                twait(ptx::smem_u32(&s_aempty[ar.i]), ar.ph ^ 1, 1);
                ptx::tc_fence_after();
            }
            const uint32_t taddr0 = taddr_base + ar.i * A_COLS + cur_s0 * 32;
            if (do_st) {
#pragma unroll
                for (int c = 0; c < CPH; c++)
                    if (h * CPH + c < 2 * cur_ns) ptx::tmem_st_32x32b_x16(taddr0 + c * 16, r[c]);
            }
            if (last) {
                if (do_wait) ptx::tmem_st_wait();
                ptx::tc_fence_before();
                __syncwarp();
                // A-ready goes to the leader; it also tells the leader that this CTA's activation half has landed
                if (lane == 0) ptx::mbar_arrive_cluster(afull_leader0 + 8u * (uint32_t) ar.i);
                ar.advance();
            }
            if (w.valid()) fetch_unpack();  // next stage: fetched and unpacked under the MMAs of this block
        }
        }   // generic decode loop
        if (q == 0 && lane == 0) tflush(2, h * 3);
    } else if (warp < 4 || (EG == 2 && warp >= 15)) {
        // ===================== epilogue: this CTA's 128 rows x NT tokens =====================
        // EG warpgroups; group eg drains the columns [eg NT / EG, (eg + 1) NT / EG) of every accumulator.  A warp reads the
        // TMEM lane quadrant warp % 4.
        ev_role = 3;
        const int eg = (EG == 1 || warp < 4) ? 0 : 1;   // (a compile-time 0 with one group: the column loops keep constant bounds)
        const int ew = (EG == 1) ? warp : (warp & 3);   // (one group: exactly the round-1 expressions -- with `warp & 3` and
                                                       //  et = eg * 128 + m the 256-token kernels spilled 320 bytes and their
                                                       //  split-tile fix-up took 30 us longer)
        const int m = ew * 32 + lane;
        const int et = (EG == 1) ? (int) threadIdx.x : eg * 128 + m;  // 0 .. ET - 1
        constexpr int CW = NT / EG;   // columns per group
        const int cbeg = eg * CW, cend = cbeg + CW;
        if (p.pdl) ptx::pdl_wait();  // reads predecessor outputs; owns all global writes of this kernel
        // staged scatter (OUT == 3): the buffers sit behind the stage ring
        epi::Scatter sct;
        sct.init(smem_base + (uint32_t) S * p.stage_bytes, et);
        epi::Scatter *const scp = STAGED ? &sct : nullptr;
        uint32_t ndrain = 0;
        const uint32_t accempty_leader0 = ptx::mapa(ptx::smem_u32(&s_accempty[0]), 0);
        const uint32_t accempty_leader1 = ptx::mapa(ptx::smem_u32(&s_accempty[1]), 0);
        // exact-path running totals: this CTA's scratch tile in the (L2-resident) workspace
        int32_t *const tots = p.g.part + ((int64_t) gridDim.x * 2 + blockIdx.x) * (NT * 128);
        // shared fix-up of split tiles (below): only when every pair owns at most ONE split unit -- all contributors of a tile
        // then arrive together, at the start of the kernel, and nobody's wait can hold up a run that another tile waits for
        // (with longer split ranges pair b's second run feeds the tile pair b + 1 waits on: a chain over all pairs)
        // (not in push mode: its copy kernel shares the SMs and waits for this kernel's stamps -- nothing here may wait on a CTA
        // that might not be resident yet)
        const bool coop = (shwhole == 2) && !p.split_last && p.coop_fix && SU < npairs && p.g.tile_done == nullptr;
        int pend_old = -1;   // thread 0: the counter of this pair's split tile before its share was counted (coop)
        Walk w;
        w.init(p, rg, shG, shgsub);
        while (w.valid()) {
            // ---- one tile run: units [ra, rb] of one output tile ----
            const int ra = w.u, rb = w.run_last();
            const int g_begin = w.g, g_end = w.g + (rb - ra);
            const bool fast = tile_fast(w.tt);
            const bool complete = (g_begin == 0) && (g_end == shG - 1);
            const int tile = w.tile;
            const int64_t n = ((int64_t) 2 * w.rp + rank) * TILE_ROWS + m;
            const int64_t t0 = (int64_t) w.tt * NT;
            if (STAGED) {
                const int64_t rows = p.g.N - (n - m);   // rows of this strip that exist
                sct.row_bytes = rows >= TILE_ROWS ? TILE_ROWS * 4 : (rows > 0 ? (int) rows * 4 : 0);
                sct.tmap = &p.out_map;
                sct.n0 = (int) (n - m);
                sct.tok = (int) t0;
            }
            // Split tile: every contributor red.adds its int32 partial into ONE accumulation tile (resolved in L2, fire
            // and forget), indexed by the first pair that owns a unit of the tile (unique among split tiles: a pair that
            // owned the first unit of two tiles would own all of the first).  The last arriver reads the sum back with
            // 64 loads in flight per thread and leaves zeros behind.  (Private slots summed by the last arriver made it
            // walk contributors x NT/16 dependent L2 round trips of ~2000 clk each under the main loop's traffic.)
            int32_t *slot = nullptr;
            if (!complete) {
                const long long x = (long long) tile * shG - sbase;
                const int b_first = (int) (((x + 1) * npairs - 1) / SU);
                slot = p.g.acc + (int64_t) (2 * b_first + rank) * (128 * NT);
            }
            for (int i = et; i < NT; i += ET) s_sc[i] = __ldg(p.g.act.scales + t0 + i);

            auto emit = [&](int c0, const int32_t (&vals)[16]) {
                if (complete) {
                    if (n < p.g.N && !(p.abl & 8)) {
#pragma unroll
                        for (int i = 0; i < 16; i++) {
                            if (t0 + c0 + i < p.g.T) {
                                float v = (float) vals[i] * s_sc[c0 + i];
                                if (p.g.w_scale != 1.0f) v *= p.g.w_scale;
                                store_out<SO>(out0, p.g.outs, p.g.n_out, (t0 + c0 + i) * p.g.ldo + n, v);
                            }
                        }
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 16; i++) atomicAdd(slot + (c0 + i) * 128 + m, vals[i]);  // RED
                }
            };

            for (int u = ra; u <= rb; u++, w.next()) {
                if (fast && u != rb) continue;  // fast path: a single drain at the end of the run
                const uint32_t ab = ndrain % NACC;
                if (flat) {
                    // the +1 code-offset correction is linear in K: the run that holds the tile's first block subtracts all
                    // of it (the K-group prefix sums do not line up with the A blocks), the other runs nothing
                    for (int i = et; i < NT; i += ET)
                        s_gs[i] = (g_begin == 0) ? __ldg(p.g.act.gpre + (t0 + i) * (p.G + 1) + p.G) -
                                                       __ldg(p.g.act.gpre + (t0 + i) * (p.G + 1))
                                                 : 0;
                } else
                for (int i = et; i < NT; i += ET)
                    s_gs[i] = __ldg(p.g.act.gpre + (t0 + i) * (p.G + 1) + w.g + 1) -
                              __ldg(p.g.act.gpre + (t0 + i) * (p.G + 1) + (fast ? g_begin : w.g));
                ptx::bar_sync(1, ET);
                twait(ptx::smem_u32(&s_accfull[ab]), (ndrain / NACC) & 1, 0);
                mark(0);
                ptx::tc_fence_after();
                const uint32_t ta = tmem_base + ((uint32_t) (ew * 32) << 16) + ab * NT;
                if (fast) {
                    // hot path: one drain per tile run, branch-free per value (epilogue.cuh)
                    const int64_t idx0 = t0 * p.g.ldo + n;
                    const int tvalid = (int) min((int64_t) NT, p.g.T - t0);
                    if (p.abl & 16) {
                        // timing ablation: the accumulator is handed back undrained
                    } else if (complete)
                        epi::drain_fast<NT, OUT, 0>(ta, s_gs, s_sc, p.g.w_scale, out0, p.g.outs, p.g.n_out, idx0, p.g.ldo, tvalid,
                                                    n < p.g.N && !(p.abl & 8) /* timing ablation: a drain without its stores */,
                                                    nullptr, p.g.vec_ok != 0, scp, cbeg, cend);
                    else
                        epi::drain_fast<NT, OUT, 2>(ta, s_gs, s_sc, p.g.w_scale, out0, p.g.outs, p.g.n_out, idx0,
                                                       p.g.ldo, tvalid, n < p.g.N, slot + m, false, nullptr, cbeg, cend);
                } else {
                    // exact path: int16 wrap per K-group, running totals in the scratch tile
#pragma unroll 1
                    for (int c0 = cbeg; c0 < cend; c0 += 16) {
                        uint32_t v[16];
                        ptx::tmem_ld_32x32b_x16(ta + c0, v);
                        ptx::tmem_ld_wait();
                        int32_t vals[16];
#pragma unroll
                        for (int i = 0; i < 16; i++) {
                            const int32_t x = (int32_t) v[i] - s_gs[c0 + i];
                            vals[i] = (u == ra) ? wrap16(x) : tots[(c0 + i) * 128 + m] + wrap16(x);
                        }
                        if (u == rb) {
                            emit(c0, vals);
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; i++) tots[(c0 + i) * 128 + m] = vals[i];
                        }
                    }
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive_cluster(ab ? accempty_leader1 : accempty_leader0);
                mark(7);
                ++ndrain;
                ptx::bar_sync(1, ET);  // s_gs is rewritten by the next drain
            }

            if (!complete) {
                // Split tile: partial parked above.  Who turns the accumulated tile into output?
                //  * stream-K: the last CTA to arrive reads the sum back and finishes the tile (nobody ever waits);
                //  * hybrid with at most one split unit per pair (split units are walked FIRST, so a tile's contributors arrive
                //    together): every contributor waits until all units are in and finishes ITS share of the 16-column chunks.
                //    The fix-up is ~6000 clk on one SM (64 KB read back over two L2 round trips, 64 KB of zeros and 64 KB of
                //    output through ~29 B/clk of store bandwidth, scripts/ubench/red_rate.cu) and made its CTA the kernel's
                //    straggler (scripts/role_life_spread.py: 44.6k clk against a mean of 36.7k at T = 512); shared by n
                //    contributors it is one round trip and 128/n KB of stores each.  The wait is bounded (below).
                mark(10);
                __threadfence();
                ptx::bar_sync(1, ET);
                int32_t *const counter = p.g.counters + tile * 2 + rank;
                const int ndone = g_end - g_begin + 1;
                if (coop) {
                    // fewer split units than pairs: every unit of the tile has its own pair, contributor g finishes the
                    // chunks g, g + G, ...  Every step below that waits for an L2 round trip (~1000-1500 clk under load) is
                    // on this pair's critical path, so there are as few as the protocol allows: fence (the partials have
                    // landed) -> add the unit count -> poll -> load the share.  The zeros and the output are fire and forget,
                    // and the count of finished shares (upper half of the counter; whoever completes it resets the counter) is
                    // only looked at when the kernel ends (pend_old): nothing orders it against the output stores -- the
                    // accumulation tile and the counter are next read by another launch.
                    // The wait needs every contributor to be running.  That holds for a grid that owns the GPU (launch_out clamps
                    // it to the occupancy), but a kernel of another stream or process may hold the SMs this grid's last CTAs are
                    // waiting for, and wait in the same way for its own -- so the wait is bounded (p.coop_timeout clk): a
                    // contributor that runs out of patience sets ABANDON (only while the tile is still incomplete: CAS) and
                    // leaves, everybody who sees the flag leaves too, and the last arriver -- which finds the flag in the value
                    // its own arrival returns -- finishes the whole tile as in the stream-K protocol.  The flag cannot change
                    // once the tile is complete, so all contributors agree on who stores what.
                    constexpr int ABANDON = 1 << 30;
                    if (et == 0) {
                        int mode = 1;   // 1: my share, 2: the whole tile, 0: nothing
                        int v = atomicAdd(counter, ndone) + ndone;
                        if ((v & 0xffff) >= shG) {
                            if (v & ABANDON) mode = 2;
                        } else if (v & ABANDON) {
                            mode = 0;
                        } else {
                            const long long w0 = clock64();
                            for (;;) {
                                __nanosleep(20);
                                asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
                                if (v & ABANDON) { mode = 0; break; }
                                if ((v & 0xffff) >= shG) break;
                                if (clock64() - w0 > (long long) p.coop_timeout && atomicCAS(counter, v, v | ABANDON) == v) {
                                    mode = 0;
                                    break;
                                }
                            }
                        }
                        __threadfence();   // acquire side for the whole epilogue (the bar.sync below carries it to the others)
                        s_last = mode;
                    }
                    ptx::bar_sync(1, ET);
                    mark(11);
                    const int mode = s_last;
                    const int cs = (mode == 1) ? cbeg + 16 * g_begin : (mode == 2 ? cbeg : cend);
                    const int cstep = (mode == 1) ? 16 * shG : 16;
#pragma unroll 1
                    for (int c0 = cs; c0 < cend; c0 += cstep) {
                        int32_t a[16];
#pragma unroll
                        for (int i = 0; i < 16; i++) a[i] = __ldcg(slot + (c0 + i) * 128 + m);
#pragma unroll
                        for (int i = 0; i < 16; i++) __stcg(slot + (c0 + i) * 128 + m, 0);
                        epi::store16<OUT>(a, s_sc + c0, p.g.w_scale != 1.0f, p.g.w_scale, out0, p.g.outs, p.g.n_out,
                                          (t0 + c0) * p.g.ldo + n, p.g.ldo, (int) min((int64_t) 16, p.g.T - t0 - c0), n < p.g.N,
                                          p.g.vec_ok != 0, scp);
                    }
                    // every contributor counts itself out exactly once (bits 16..29); the last one leaves the counter at zero
                    if (et == 0) pend_old = atomicAdd(counter, 1 << 16);   // not waited for here
                    mark(12);
                } else {
                if (et == 0) {
                    const int old = atomicAdd(counter, ndone);
                    s_last = (old + ndone == shG);
                }
                ptx::bar_sync(1, ET);
                mark(11);
                if (s_last) {
                    __threadfence();
                    constexpr int CB = (CW >= 64) ? 4 : CW / 16;  // 16-column chunks per round trip
#pragma unroll 1
                    for (int c0 = cbeg; c0 < cend; c0 += 16 * CB) {
                        int32_t a[CB][16];
#pragma unroll
                        for (int j = 0; j < CB; j++)
#pragma unroll
                            for (int i = 0; i < 16; i++) a[j][i] = __ldcg(slot + (c0 + 16 * j + i) * 128 + m);
#pragma unroll
                        for (int j = 0; j < CB; j++)
#pragma unroll
                            for (int i = 0; i < 16; i++) __stcg(slot + (c0 + 16 * j + i) * 128 + m, 0);
#pragma unroll
                        for (int j = 0; j < CB; j++)
                            epi::store16<OUT>(a[j], s_sc + c0 + 16 * j, p.g.w_scale != 1.0f, p.g.w_scale, out0, p.g.outs,
                                              p.g.n_out, (t0 + c0 + 16 * j) * p.g.ldo + n, p.g.ldo,
                                              (int) min((int64_t) 16, p.g.T - t0 - c0 - 16 * j), n < p.g.N, p.g.vec_ok != 0, scp);
                    }
                    if (et == 0) *counter = 0;
                    mark(12);
                }
                }
            }
            // push mode: this CTA's block of the tile is in the local result buffer -> stamp it for the copy kernel
            const bool stamp = p.g.tile_done != nullptr && (complete || s_last);
            if (stamp) __threadfence();
            ptx::bar_sync(1, ET);  // s_sc / s_last are reused by the next run
            if (stamp && et == 0)
                asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.g.tile_done + tile * 2 + rank), "r"(p.g.tile_epoch) : "memory");
        }
        if (et == 0 && pend_old >= 0 && ((pend_old >> 16) & 0x3fff) + 1 == shG)   // last contributor out: leave the counter at zero
            p.g.counters[(rg.b0 / shG) * 2 + rank] = 0;
        if (et == 0) tflush(3, 0);
        if (STAGED) sct.finish();   // every bulk copy of this thread has been written out before the CTA may exit
    }

    // ---- teardown: nobody leaves (or frees TMEM) while the peer may still read this CTA's shared memory / TMEM ----
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    if (warp == 14) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_pair(tmem_base, 512);
    }
}

template <int NT, int SPS, int OUT, int SPA, int NA>
static cudaError_t launch_out(Params &p, int sm_count, cudaStream_t s) {
    const Knobs &kn = knobs();
    // 200 KB by default; the 256-token tile (40 KB stages) takes 5 stages = 201 KB + alignment slack
    const int budget = (NT > 128 && kn.tc_smem_kb == 200 ? 204 : kn.tc_smem_kb) * 1024;
    p.stage_bytes = SPS * (NT / 2 * 128 + SUB_W_BYTES);  // multiple of 1024
    int ns = (budget - 1024) / p.stage_bytes;
    {   // the staged drain's buffers share the 227 KB with the ring (and ~3 KB of static shared memory)
        constexpr int scatter = (OUT == 3 || OUT == 4) ? epi::SCATTER_BYTES : 0;
        const int fit = (224 * 1024 - 1024 - scatter) / p.stage_bytes;
        if (ns > fit) ns = fit;
    }
    if (ns > MAX_STAGES) ns = MAX_STAGES;
    int forced = kn.tc_stages;
    // the decode warps fetch the first stage of the next A block before they hand over the current one: the ring must
    // hold one stage more than an A block, or that fetch would wait for an MMA that waits for the hand-over
    constexpr int MIN_STAGES = SPA / SPS + 1;
    if (forced > 0 && forced < MIN_STAGES) forced = MIN_STAGES;
    if (forced > 0 && forced < ns) ns = forced;
    if (ns < MIN_STAGES) return cudaErrorInvalidValue;
    p.nstages = ns;
    const size_t smem = (size_t) ns * p.stage_bytes + 1024 + ((OUT == 3 || OUT == 4) ? epi::SCATTER_BYTES : 0);
    auto kern = kernel<NT, SPS, OUT, SPA, NA>;
    static int smem_set[MAX_DEVICES] = {};
    cudaError_t e = ensure_dyn_smem(kern, smem, p.g.device, smem_set);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(threads_for(OUT));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = p.pdl ? 1 : 0;
    // persistent: one CTA pair per co-resident cluster
    int npairs = sm_count / 2;
    static int cached = 0;   // per (NT, SPS, OUT) instantiation
    if (cached == 0) {
        cfg.gridDim = dim3((unsigned) (2 * npairs));
        int n = 0;
        e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
        if (e != cudaSuccess) return e;
        if (n < 1) return cudaErrorInvalidConfiguration;
        cached = n;
    }
    if (npairs > cached) npairs = cached;
    // Work split, in K-group units of pair time:
    //   whole (1)   rounds * G                         whole tiles only; pairs beyond ptiles / rounds stay idle
    //   hybrid (2)  full * G + 2 ceil(SU / npairs) + 3  floor(ptiles / npairs) whole tiles per pair; the SU units of the
    //                                                  remaining tiles are split over all pairs and done FIRST, so the
    //                                                  fix-up (slot write, fence, counter, the last arriver's reduction)
    //                                                  runs in the epilogue warps under the MMAs of the whole tiles
    //   stream-K (0) ceil(U / npairs) + 10             every pair ends on a split tile: the fix-up is exposed at the tail
    //                                                  (measured T=512: 39.6 us vs 33.9 us whole on 56 of the 74 pairs)
    const int ptiles = p.RP * p.TT;
    const int rounds = (ptiles + npairs - 1) / npairs;
    const int forced_grid = kn.tc_grid;  // tests: force a grid (and with it the split paths)
    auto policy = [&](int G, int U, int gsub) {    // G units per tile, U in all, gsub 128-k sub-blocks per unit
        // costs in 512-k units of pair time (the fix-up constants were measured with 512-k K-groups)
        const double scale = gsub / 4.0;
        const int full = ptiles / npairs, su = (ptiles - full * npairs) * G;
        const double cost_whole = rounds * G * scale;
        // refitted on the lean-loop kernel (scripts/time_split.py, profiles/r2_tcp_split_policies.txt): a split unit costs
        // ~1.2 units; the fix-up of a split tile costs ~7 units next to short whole tiles (G = 8) and hides almost
        // completely under long ones (G = 28); stream-K's exposed fix-up ~16
        const double tile_units = G * scale;
        // (... and a little better the more whole tiles follow it: with three or more, hybrid wins the ties -- all SMs busy)
        const double hybrid_fix = (tile_units > 24.0 ? 2.0 : 10.0 - tile_units / 3.0) * (full >= 3 ? 0.85 : 1.0);
        const double cost_hybrid = full * tile_units + (su ? 1.2 * ((su + npairs - 1) / npairs) * scale + hybrid_fix : 0);
        const double cost_sk = ((U + npairs - 1) / npairs) * scale + 16.0;
        int whole = 1;
        if (full >= 1 && cost_hybrid < cost_whole && cost_hybrid <= cost_sk) whole = 2;
        else if (cost_sk < cost_whole) whole = 0;
        if (forced_grid > 0) whole = 0;
        if (kn.tc_whole >= 0) whole = kn.tc_whole;
        return whole;
    };
    // flat walk (see Params): for shapes the lean loops cannot take as they are, when a block is one shared-memory stage
    p.uniform_knob = kn.tc_lean;
    p.flat_ok = (kn.tc_flat && kn.tc_lean && p.allow_fast && SPS == SPA && !(p.uniform && p.gsub % SPA == 0)) ? 1 : 0;
    p.Gf = (p.KB + SPA - 1) / SPA;
    p.Uf = p.RP * p.TT * p.Gf;
    p.whole = policy(p.G, p.U, p.gsub);
    p.wholef = policy(p.Gf, p.Uf, SPA);
    p.split_last = kn.tc_split_last;
    p.coop_fix = kn.tc_coop;
    p.coop_timeout = kn.tc_coop_timeout;
    if (forced_grid > 0) npairs = forced_grid / 2;
    else if (p.whole == 1 && !p.flat_ok) npairs = (ptiles + rounds - 1) / rounds;  // the fewest pairs that finish in the same number of rounds
    if (npairs > p.U) npairs = p.U;
    if (p.flat_ok && npairs > p.Uf) npairs = p.Uf;
    if (2 * npairs > sm_count) npairs = sm_count / 2;  // one accumulation tile per CTA of a resident grid (GemmArgs.acc)
    if (npairs < 1) npairs = 1;
    cfg.gridDim = dim3((unsigned) (2 * npairs));
    return cudaLaunchKernelEx(&cfg, kern, p);
}

template <int NT, int SPS, int SPA, int NA>
static cudaError_t launch_one(Params &p, int sm_count, cudaStream_t s) {
    if (p.g.n_out == 0) return p.g.local_bulk ? launch_out<NT, SPS, 4, SPA, NA>(p, sm_count, s) : launch_out<NT, SPS, 0, SPA, NA>(p, sm_count, s);
    if (p.g.multicast) return launch_out<NT, SPS, 2, SPA, NA>(p, sm_count, s);
    // peer pointers: staged bulk-copy scatter when the destinations allow it (16-byte aligned rows, N % 4 == 0)
    return p.g.bulk_ok ? launch_out<NT, SPS, 3, SPA, NA>(p, sm_count, s) : launch_out<NT, SPS, 1, SPA, NA>(p, sm_count, s);
}


// one (token tile, stage layout) configuration, every epilogue flavour
#define VLUT_TC5_INSTANCE(NAME, NT, SPS, SPA, NA) \
    cudaError_t NAME(Params &p, int sm_count, cudaStream_t s) { return launch_one<NT, SPS, SPA, NA>(p, sm_count, s); }

}  // namespace tc5
}  // namespace vlut
