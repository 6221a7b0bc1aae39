// gemm_tcs.cu -- VLUT_ALGO_TCS: the small-token-count Vec-LUT GEMM (T <= 64), built to STREAM the packed weights from HBM.
//
// Same arithmetic and exactness contract as gemm_tc4.cu (tcgen05 kind::i8, u8 weight codes x s8 activations -> s32 in TMEM,
// per-K-group int16 wrap on the exact path, fused fp32 epilogue), reorganised around what in-kernel
// timelines showed at T <= 64 (scripts/timeline_tcs.py, scripts/ubench/umma_contend.cu): nothing is bandwidth- or
// throughput-bound there; every role is ONE warp walking a serial chain of waits and dependent scalar instructions
// (~600 clk per 256-k block for a decode warp, ~50 clk of issue per MMA for the MMA thread), so the kernel is built from
// many short independent chains instead of a few long ones.
//
//   * WEIGHTS never pass through a producer warp or an mbarrier.  Decode thread (q, h, p, lane) owns weight row
//     m = 32q + lane of the strip, sub-block h of the 256-k A blocks of parity p; it fetches its own 2 x 16 bytes per
//     block with cp.async (LDGSTS, coalesced 512 B per warp instruction, L2 evict_first) into a private ring of DT blocks
//     in shared memory and reads them back itself (cp.async.wait_group: no cross-thread synchronisation at all).
//     DT x 16 KB per CTA are in flight from the first instruction of the kernel -- before griddepcontrol.wait, i.e. while
//     the quantise kernel is still running.
//   * SIXTEEN decode warps (two parities) and TWO MMA issuer warps (even / odd A blocks, one TMEM accumulator each, summed
//     by the epilogue): four A stages in TMEM, two independent decode -> MMA chains.
//   * K is SPLIT over the CTAs of a thread-block cluster (S = 1..8 CTAs per 128-row strip, whole K-groups each) so that
//     shapes with few strips (4096 -> 4096: 32) still put >= 128 CTAs on the weight stream.  Partial sums meet in
//     distributed shared memory: every CTA parks its int32 tile in its own shared memory, one cluster barrier, then CTA r
//     reduces and stores the token slice r of the tile (ld.shared::cluster).  Integer partials: order-independent, exact.
//
//   warps 0-3   epilogue (TMEM lane quadrant = warp)
//   warps 4-19  decode: d = warp - 4, quadrant q = d % 4, sub-block h = (d / 4) % 2, block parity p = d / 8
//   warp 20     TMEM allocator + activation producer (TMA)     warps 21, 22  MMA issuers (even / odd blocks)
#include "epilogue.cuh"
#include "ptx.cuh"
#include "vlut_common.cuh"

#ifndef VLUT_TCS_DEC_MARKS
#define VLUT_TCS_DEC_MARKS 15
#endif
#define DECM(bit) ((VLUT_TCS_DEC_MARKS & (bit)) != 0)
#ifndef VLUT_TCS_ABL
#define VLUT_TCS_ABL 0   // timing / bisection builds (wrong results): 1 no tcgen05.st, 2 no tcgen05.mma, 4 no tcgen05.ld, 8 no cp.async,
                         // 16 no act copy, 32 no wait::st, 64 decode ignores A-stage-free, 128 MMA ignores A-stage-full
#endif
#define ABL(bit) ((VLUT_TCS_ABL & (bit)) != 0)
#ifndef VLUT_TCS_BACKOFF
#define VLUT_TCS_BACKOFF 0   // which waits sleep between polls: 1 epilogue, 2 activation producer, 4 MMA issuers, 8 decode
#endif
#define WAIT(bit, ns, bar, par) do { if (VLUT_TCS_BACKOFF & (bit)) ptx::mbar_wait_backoff(bar, par, ns); else ptx::mbar_wait(bar, par); } while (0)

namespace vlut {

namespace tcs {

constexpr int W_DEC0 = 4, NDEC = 16, W_ACT = 20, W_MMA0 = 21;
constexpr int THREADS = 23 * 32;
constexpr int SPA = 2;               // 128-k sub-blocks per A block (decode -> MMA hand-off): 256 k
constexpr int A_COLS = SPA * 32;     // TMEM columns of one A stage
constexpr int NPH = 4;               // decode phases: the four warps of phase ph unpack the blocks b = ph (mod 4)
constexpr int TMEM_COLS = 512;
constexpr int MAX_AS = 8;            // activation stages (ring)
constexpr int DEC_THREADS = NDEC * 32;
// MASK (token tiles <= 32): mask-only weight unpack.  The 2-bit codes are not shifted down: word & (0x03030303 << 2i) is the
// A operand of bit-pair position i, scaled by 4^i, and goes to accumulator i (4 per issuer); the epilogue adds D_i >> 2i.
// The activation rows are regrouped to match (ActView.perm).  16 LOP3 per 64 weights instead of 16 LOP3 + 12 IMAD.HI.
__host__ __device__ constexpr bool mask_mode(int nt) { return nt <= 32; }
__host__ __device__ constexpr int accs_for(int nt) { return mask_mode(nt) ? 4 : 1; }                 // accumulators per issuer
__host__ __device__ constexpr int na_for(int nt) { return (TMEM_COLS - 2 * accs_for(nt) * nt) / A_COLS > 6 ? 6 : (TMEM_COLS - 2 * accs_for(nt) * nt) / A_COLS; }
constexpr int W_BLOCK_BYTES = 2 * SPA * TILE_BYTES;  // packed bytes of one A block of a strip: 8 KB

struct Params {
    GemmArgs g;
    int KB, G, gsub, RT;
    int S;          // CTAs per strip (cluster size): K-groups split [r*G/S, (r+1)*G/S)
    int R;          // weight rows per (virtual) strip, <= 128: rows [rt*R, rt*R + R) of the tensor; 128 = the T2 strips themselves
    int NAS, ASB;   // activation ring: NAS stages of ASB 128-k sub-blocks (ASB a power of two >= 4)
    int asb_shift;
    int allow_fast, pdl;
    int issuers;    // MMA issuer warps used on the fast path (2; 1 = experiment / fallback)
    int pre;        // weight-ring slots requested before the activation tile is on its way (the rest right after)
    long long *dbg;   // optional timeline of one CTA (VLUT_TC_DEBUG=1): [4 roles][256] x (clock64, kind)
    int dbg_cta, dbg_mask;
};

__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {  // K-major, SWIZZLE_128B, 8-row atoms 1024 B apart
    uint64_t d = 0;
    d |= (uint64_t) ((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t) 1 << 16;
    d |= (uint64_t) (1024 >> 4) << 32;
    d |= (uint64_t) 1 << 46;
    d |= (uint64_t) 2 << 61;
    return d;
}
__host__ __device__ constexpr uint32_t make_idesc(int n) {  // D s32, A u8, B s8, K-major both, M = 128, N = n
    return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (128 >> 4) << 24);
}

template <int NT, int OUT, int DT, bool DBG>
__global__ void __launch_bounds__(THREADS, 1) kernel(const __grid_constant__ Params p) {
    constexpr bool MASK = mask_mode(NT);
    constexpr int ACCS = accs_for(NT);                // accumulators per MMA issuer (fast path) / per drain set (exact path)
    constexpr int NA = na_for(NT);                    // A stages in TMEM; block b uses stage b % NA
    constexpr uint32_t A_COL0 = 2 * ACCS * NT;        // two accumulator sets, then the A stages
    static_assert(NA >= 4 && A_COL0 + NA * A_COLS <= TMEM_COLS, "TMEM budget");
    constexpr int ACT_BYTES = NT * 128;               // one 128-k sub-block of the token tile, pre-swizzled by the quantiser
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_actfull[MAX_AS], s_actempty[MAX_AS];
    __shared__ __align__(8) uint64_t s_afull[6], s_aempty[6];
    __shared__ __align__(8) uint64_t s_accfull[2], s_accempty[2];
    __shared__ uint32_t s_tmem_base;
    __shared__ __align__(16) int32_t s_gs[NT];
    __shared__ __align__(16) float s_sc[NT];

    // the warp index through a shuffle: the compiler then knows it is warp-uniform and keeps the role loops' state (block
    // counters, barrier and TMEM addresses, descriptors) in uniform registers instead of converting per use (R2UR)
    const int warp = __shfl_sync(0xffffffffu, (int) (threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int S = p.S;
    const int rank = (S > 1) ? (int) ptx::cluster_ctarank() : 0;
    // strip of the launch's concatenated strip space -> (tensor of the batch, strip inside it)
    const int rt_all = blockIdx.x / S;
    int ti = 0;
#pragma unroll
    for (int i = 1; i < MAX_BATCH; i++) ti += (i < p.g.n_tens && rt_all >= p.g.tens[i].strip0) ? 1 : 0;
    const TensorRef &ten = p.g.tens[ti];
    const int rt = rt_all - ten.strip0;
    const int tt = blockIdx.y;
    const int g0 = (rank * p.G) / S, g1 = ((rank + 1) * p.G) / S;   // this CTA's K-groups
    const int kb_begin = g0 * p.gsub, kb_end = min(g1 * p.gsub, p.KB);
    const int nsub_slice = kb_end - kb_begin;
    // A blocks of this CTA: pairs of 128-k sub-blocks over the K slice, whatever the K-groups are (group boundaries only
    // concern the MMA issuers and the epilogue)
    const int nblk = (nsub_slice + SPA - 1) / SPA;

    // dynamic shared memory: [activation ring][weight ring: DT slots x 4 chunks x 512 threads x 16 B][int32 tile NT x 128]
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t act_base = smem_base;
    const uint32_t wring = act_base + (uint32_t) (p.NAS * p.ASB) * ACT_BYTES;
    const uint32_t red_base = wring + DT * 4 * DEC_THREADS * 16;
    int32_t *const red = reinterpret_cast<int32_t *>(smem_raw + (red_base - ptx::smem_u32(smem_raw)));

    // optional timeline (absolute SM clock, so the roles of one CTA share a time base).  Compiled in only for DBG kernels;
    // branch-free (a predicated store), so that it cannot de-converge a warp in front of an aligned tcgen05 instruction.
    const uint32_t dbg_lane0 = (DBG && p.dbg != nullptr && (int) blockIdx.x == p.dbg_cta && lane == 0) ? 1u : 0u;
    int n_ev = 0, ev_role = 0;
    auto mark = [&](int kind) {
        if constexpr (DBG) {
#ifdef VLUT_TCS_MARK_ROLES
            if (!((VLUT_TCS_MARK_ROLES >> ev_role) & 1)) return;
#endif
            const uint32_t on = (dbg_lane0 && n_ev < 256 && ((p.dbg_mask >> ev_role) & 1)) ? 1u : 0u;
            long long *dst = p.dbg + (ev_role * 256 + min(n_ev, 255)) * 2;
            const long long now = clock64();
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %0, 0;\n\t@p st.global.v2.u64 [%1], {%2, %3};\n\t}" ::"r"(on),
                         "l"(dst), "l"(now), "l"((long long) kind)
                         : "memory");
            ++n_ev;
        }
    };

    // let the next kernel of the stream start its prologue as early as resources allow
    if (p.pdl) ptx::pdl_launch_dependents();

    if (warp >= W_DEC0 && warp < W_DEC0 + NDEC) {
        // ===================== decode: HBM -> (cp.async ring) -> registers -> u8 A operand in TMEM =====================
        const int d = warp - W_DEC0;
        const int q = d & 3, ph = d >> 2;
        const int m = q * 32 + lane;
        const int dt = d * 32 + lane;
        ev_role = 0;
        const bool dmark = (d == 0);
        // this thread unpacks its row of the blocks b = ph, ph + 4, ...: 4 x 16 B per block (two 128-k sub-blocks of two
        // 64-k chunks), 32 KB further on each time
        // virtual strips (R < 128 rows, so that e.g. 14336 rows give 144 CTAs on 148 SMs instead of 112): row n of the tensor
        // sits in T2 strip n / 128 at row n % 128; lanes beyond the strip's R rows (or the padded matrix) fetch nothing
        const int64_t nrow = (int64_t) rt * p.R + m;
        const bool wrow = m < p.R && nrow < (int64_t) round_up(ten.N, TILE_ROWS);
        const uint8_t *wsrc = ten.w + ((int64_t) (nrow >> 7) * (2 * p.KB) + 2 * kb_begin) * TILE_BYTES + (nrow & 127) * 16;
        const uint32_t wdst = wring + dt * 16;
        auto issue = [&](int b, int slot) {   // this thread's 64 bytes of block b -> ring slot; always one commit group
            if (!ABL(8) && b < nblk && wrow) {
                const uint8_t *src = wsrc + (int64_t) b * W_BLOCK_BYTES;
                ptx::cp_async16(wdst + (slot * 4 + 0) * (DEC_THREADS * 16), src);
                ptx::cp_async16(wdst + (slot * 4 + 1) * (DEC_THREADS * 16), src + TILE_BYTES);
                if (SPA * b + 1 < nsub_slice) {
                    ptx::cp_async16(wdst + (slot * 4 + 2) * (DEC_THREADS * 16), src + 2 * TILE_BYTES);
                    ptx::cp_async16(wdst + (slot * 4 + 3) * (DEC_THREADS * 16), src + 3 * TILE_BYTES);
                }
            }
            ptx::cp_async_commit();
        };
        // weights are static data: the stream starts now, whatever the predecessor kernel (quantise) is still doing.
        // (`pre` < DT is an experiment knob, VLUT_TCS_PRE: request only `pre` ring slots here and the rest once the
        // activation producer has its first stage on the way.  The in-kernel timeline shows the first MMA ~2200 clk after
        // the first weights land; the hypothesis that the activation copies queue behind the 10 MB of weight prefetch did
        // not hold: delaying the prefetch measured 0.3-0.6 us slower, profiles/r2_tcs_pre.txt.)
#pragma unroll
        for (int i = 0; i < DT; i++)
            if (i < p.pre) issue(ph + NPH * i, i);
        // barriers / TMEM address are set up by warp 20 meanwhile
        ptx::bar_sync(2, THREADS);
        ptx::tc_fence_after();
        if (p.pre < DT) {
            ptx::bar_sync(3, DEC_THREADS + 32);   // the activation producer arrives after issuing its first copy
#pragma unroll
            for (int i = 0; i < DT; i++)
                if (i >= p.pre) issue(ph + NPH * i, i);
        }
        const uint32_t tmem_base = s_tmem_base;
        const uint32_t taddr0 = tmem_base + ((uint32_t) (q * 32) << 16) + A_COL0;
        int slot = 0;
        int ai = ph % NA;              // A stage of block b (= b % NA) and how often it has been used before (b / NA)
        uint32_t an = ph / NA;
#pragma unroll 1
        for (int b = ph; b < nblk; b += NPH) {
            ptx::cp_async_wait<DT - 1>();
            if (dmark && DECM(1)) mark(0);
            const bool two = SPA * b + 1 < nsub_slice;
            uint4 cw[4];
#pragma unroll
            for (int c = 0; c < 4; c++) {
                if (c < 2 || two) cw[c] = ptx::lds128(wdst + (slot * 4 + c) * (DEC_THREADS * 16));
                else cw[c] = make_uint4(0x55555555u, 0x55555555u, 0x55555555u, 0x55555555u);   // weight 0 everywhere
            }
            issue(b + NPH * DT, slot);   // refill the slot just read (same thread: program order)
            // the first sub-block is unpacked into registers BEFORE the A stage is known to be free, the second one while
            // the first store is in flight (64 + 16 registers for both at once would not fit 736 threads)
            auto unpack = [&](int sb, uint32_t (&r)[32]) {
                const uint32_t wv[8] = {cw[2 * sb].x, cw[2 * sb].y, cw[2 * sb].z, cw[2 * sb].w,
                                        cw[2 * sb + 1].x, cw[2 * sb + 1].y, cw[2 * sb + 1].z, cw[2 * sb + 1].w};
                if constexpr (MASK) {
                    // column i*8 + w: the codes at bit-pair position i of word w, left in place (scaled by 4^i)
#pragma unroll
                    for (int i = 0; i < 4; i++)
#pragma unroll
                        for (int w8 = 0; w8 < 8; w8++) r[i * 8 + w8] = wv[w8] & (0x03030303u << (2 * i));
                } else {
                    // column 4w + i: shifts as multiply-high on the FMA pipe, masks on the ALU pipe
#pragma unroll
                    for (int w8 = 0; w8 < 8; w8++) {
                        r[4 * w8 + 0] = wv[w8] & 0x03030303u;
                        r[4 * w8 + 1] = __umulhi(wv[w8], 1u << 30) & 0x03030303u;
                        r[4 * w8 + 2] = __umulhi(wv[w8], 1u << 28) & 0x03030303u;
                        r[4 * w8 + 3] = __umulhi(wv[w8], 1u << 26) & 0x03030303u;
                    }
                }
            };
            uint32_t r0[32];
            unpack(0, r0);
            if (!ABL(64)) WAIT(8, 20, ptx::smem_u32(&s_aempty[ai]), (an & 1) ^ 1);
            ptx::tc_fence_after();
            if (dmark && DECM(2)) mark(1);
            if (!ABL(1)) ptx::tmem_st_32x32b_x32(taddr0 + ai * A_COLS, r0);
            if (two) {
                uint32_t r1[32];
                unpack(1, r1);
                if (!ABL(1)) ptx::tmem_st_32x32b_x32(taddr0 + ai * A_COLS + 32, r1);
            }
            if (dmark && DECM(4)) mark(2);
            if (!ABL(32)) ptx::tmem_st_wait();
            if (dmark && DECM(8)) mark(3);
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_afull[ai]));
            __syncwarp();
            if (++slot == DT) slot = 0;
            ai += NPH % NA;
            an += NPH / NA;
            if (ai >= NA) { ai -= NA; ++an; }
        }
        ptx::cp_async_wait<0>();
    } else {
        if (warp == W_ACT) {
            if (lane == 0) {
                for (int s = 0; s < p.NAS; s++) {
                    ptx::mbar_init(ptx::smem_u32(&s_actfull[s]), 1);
                    ptx::mbar_init(ptx::smem_u32(&s_actempty[s]), 2);   // one commit per MMA issuer
                }
                for (int d = 0; d < NA; d++) {
                    ptx::mbar_init(ptx::smem_u32(&s_afull[d]), NDEC / NPH);   // the four decode warps of the block's phase
                    ptx::mbar_init(ptx::smem_u32(&s_aempty[d]), 1);
                }
                for (int d = 0; d < 2; d++) {
                    ptx::mbar_init(ptx::smem_u32(&s_accfull[d]), 1);
                    ptx::mbar_init(ptx::smem_u32(&s_accempty[d]), 4);
                }
                ptx::fence_barrier_init();
            }
            __syncwarp();
            ptx::tmem_alloc(ptx::smem_u32(&s_tmem_base), TMEM_COLS);
            ptx::tmem_relinquish();
            ptx::tc_fence_before();
        }
        ptx::bar_sync(2, THREADS);
        ptx::tc_fence_after();
        const uint32_t tmem_base = s_tmem_base;
        // everything below reads what the predecessor kernel (quantise) wrote, or writes global memory
        if (p.pdl) ptx::pdl_wait();
        const bool fast = p.allow_fast && (__ldg(p.g.act.tile_unsafe + tt) != p.g.act.epoch);

        if (warp == W_ACT) {
            // ===================== activation producer: one bulk copy per stage of ASB sub-blocks =====================
            ev_role = 3;
            mark(0);
            const uint64_t pol = ptx::policy_evict_last();
            const int n_as = (nsub_slice + p.ASB - 1) >> p.asb_shift;
            const int8_t *src0 = p.g.act.q_sw + ((int64_t) tt * p.KB + kb_begin) * ACT_BYTES;
            for (int asi = 0, ring = 0, ph = 0; asi < n_as; asi++) {
                WAIT(2, 100, ptx::smem_u32(&s_actempty[ring]), ph ^ 1);
                if (ptx::elect_one()) {
                    const uint32_t bytes = (uint32_t) min(p.ASB, nsub_slice - asi * p.ASB) * ACT_BYTES;
                    const uint32_t full = ptx::smem_u32(&s_actfull[ring]);
                    if (ABL(16)) ptx::mbar_arrive(full);
                    else ptx::mbar_arrive_expect_tx(full, bytes);
                    if (!ABL(16)) ptx::bulk_g2s_hint(act_base + (uint32_t) (ring * p.ASB) * ACT_BYTES, src0 + (int64_t) asi * p.ASB * ACT_BYTES,
                                       bytes, full, pol);
                }
                __syncwarp();
                mark(1);
                if (asi == 0 && p.pre < DT) ptx::bar_arrive(3, DEC_THREADS + 32);   // releases the rest of the weight prefetch
                if (++ring == p.NAS) { ring = 0; ph ^= 1; }
            }
            if (n_as == 0 && p.pre < DT) ptx::bar_arrive(3, DEC_THREADS + 32);
        } else if (warp >= W_MMA0) {
            // ===================== MMA issuers =====================
            constexpr uint32_t idesc = make_idesc(NT);
            constexpr uint32_t SUBD = ACT_BYTES >> 4;   // shared-memory descriptor units per 128-k sub-block
            const int w = warp - W_MMA0;
            ev_role = 1;
            if (w == 0) mark(10);
            const uint64_t desc0 = make_smem_desc(act_base);
            // One K=32 MMA per (sub-block, quarter ks): A columns ks*8 .. +8 of the sub-block, B bytes ks*32 .. +32 of the
            // activation rows.  MASK: quarter ks is bit-pair position ks of the weight words -> accumulator ks.
            if (fast) {
                // Hot loop.  One thread feeds the tensor pipe and its scalar instruction stream is a critical path at N = 16
                // (scripts/ubench/umma_contend.cu: 40-65 clk per MMA with per-MMA operand set-up and a re-election per
                // sub-block, against ~18 clk for the MMA itself): per A block one wait, one election and one straight-line
                // burst of 8 MMAs + commits, every operand an increment of a uniform register -- and two issuers, on
                // alternate blocks, each with its own accumulator set (the epilogue adds the two).
                const uint32_t d_tmem = tmem_base + w * (ACCS * NT);
                int asi_seen = -1, ring = 0;
                uint32_t as_ph = 0;
                const int bstep = p.issuers;
                const int b0 = (bstep == 2 ? w : (w == 0 ? 0 : nblk));
                int ai = b0 % NA;
                uint32_t an = b0 / NA;
#pragma unroll 1
                for (int b = b0; b < nblk; b += bstep) {
                    const int rel = SPA * b, asi = rel >> p.asb_shift, asr = rel & (p.ASB - 1);
                    if (!ABL(128)) WAIT(4, 20, ptx::smem_u32(&s_afull[ai]), an & 1);
                    if (asi != asi_seen) {
                        // stages are visited in order, one or two steps at a time (a stage holds >= 2 blocks)
                        for (; asi_seen < asi; ++asi_seen)
                            if (asi_seen >= 0 && ++ring == p.NAS) { ring = 0; as_ph ^= 1; }
                        ptx::mbar_wait(ptx::smem_u32(&s_actfull[ring]), as_ph);
                    }
                    ptx::tc_fence_after();
                    if (w == 0) mark(0);
                    const bool two = rel + 1 < nsub_slice;
                    const bool first = (b == b0);
                    const bool last = b + bstep >= nblk;
                    const bool rel_as = last || ((SPA * (b + bstep)) >> p.asb_shift) != asi;   // my last block in this stage
                    if (ptx::elect_one()) {
                        const uint32_t a_tmem = tmem_base + A_COL0 + ai * A_COLS;
                        const uint64_t bd = desc0 + (uint64_t) ((uint32_t) (ring * p.ASB + asr) * SUBD);
#pragma unroll
                        for (int ks = 0; ks < 4; ks++)
                            if (!ABL(2))
                                ptx::umma_i8_ts(d_tmem + (MASK ? ks * NT : 0), a_tmem + ks * 8, bd + (uint64_t) (ks * 2), idesc,
                                                (first && (MASK || ks == 0)) ? 0u : 1u);
                        if (two) {
#pragma unroll
                            for (int ks = 0; ks < 4; ks++)
                                if (!ABL(2))
                                    ptx::umma_i8_ts(d_tmem + (MASK ? ks * NT : 0), a_tmem + 32 + ks * 8, bd + (uint64_t) (SUBD + ks * 2),
                                                    idesc, 1u);
                        }
                        ptx::umma_commit(ptx::smem_u32(&s_aempty[ai]));
                        if (rel_as) {
                            ptx::umma_commit(ptx::smem_u32(&s_actempty[ring]));
                            if (bstep == 1) ptx::umma_commit(ptx::smem_u32(&s_actempty[ring]));   // both arrivals are mine
                        }
                        if (last) ptx::umma_commit(ptx::smem_u32(&s_accfull[w]));
                    }
                    __syncwarp();
                    if (w == 0) mark(1);
                    ai += bstep;
                    if (ai >= NA) { ai -= NA; ++an; }
                }
                if (b0 >= nblk) {   // no block at all for this issuer: nothing to drain, nothing to wait for
                    if (ptx::elect_one()) ptx::umma_commit(ptx::smem_u32(&s_accfull[w]));
                    __syncwarp();
                }
            } else if (w == 0) {
                // Exact path (some token of the tile can overflow int16 inside a K-group): one accumulator drain per K-group,
                // the two accumulator sets alternate.  One issuer, generic per-sub-block loop; group boundaries may fall
                // inside an A block.
                uint32_t ndrain = 0, ab = 0;
                int g = g0, gpos = 0;          // current K-group and sub-block position inside it
                int ring = 0, asr = 0;
                uint32_t as_ph = 0;
                bool fresh = true;
                int ai = 0;
                uint32_t an = 0;
#pragma unroll 1
                for (int b = 0; b < nblk; b++) {
                    ptx::mbar_wait(ptx::smem_u32(&s_afull[ai]), an & 1);
                    ptx::tc_fence_after();
                    const int ns = min(SPA, nsub_slice - SPA * b);
                    for (int sb = 0; sb < ns; sb++) {
                        const int kb = kb_begin + SPA * b + sb;
                        if (gpos == 0) {
                            ab = ndrain & 1;
                            ptx::mbar_wait(ptx::smem_u32(&s_accempty[ab]), ((ndrain >> 1) & 1) ^ 1);
                            ptx::tc_fence_after();
                            fresh = true;
                        }
                        if (asr == 0) ptx::mbar_wait(ptx::smem_u32(&s_actfull[ring]), as_ph);
                        const bool rel_as = (asr == p.ASB - 1) || (kb == kb_end - 1);
                        const bool group_end = (gpos + 1 == min(p.gsub, p.KB - g * p.gsub));
                        if (ptx::elect_one()) {
                            const uint32_t d_tmem = tmem_base + ab * (ACCS * NT);
                            const uint32_t a_tmem = tmem_base + A_COL0 + ai * A_COLS + sb * 32;
                            const uint64_t bd = desc0 + (uint64_t) ((uint32_t) (ring * p.ASB + asr) * SUBD);
#pragma unroll
                            for (int ks = 0; ks < 4; ks++)
                                ptx::umma_i8_ts(d_tmem + (MASK ? ks * NT : 0), a_tmem + ks * 8, bd + (uint64_t) (ks * 2), idesc,
                                                (fresh && (MASK || ks == 0)) ? 0u : 1u);
                            if (rel_as) {   // both arrivals of the stage's release barrier come from this issuer here
                                ptx::umma_commit(ptx::smem_u32(&s_actempty[ring]));
                                ptx::umma_commit(ptx::smem_u32(&s_actempty[ring]));
                            }
                            if (group_end) ptx::umma_commit(ptx::smem_u32(&s_accfull[ab]));
                        }
                        __syncwarp();
                        fresh = false;
                        if (group_end) { ++ndrain; ++g; gpos = 0; } else { ++gpos; }
                        if (rel_as) {
                            asr = 0;
                            if (++ring == p.NAS) { ring = 0; as_ph ^= 1; }
                        } else {
                            ++asr;
                        }
                    }
                    if (ptx::elect_one()) ptx::umma_commit(ptx::smem_u32(&s_aempty[ai]));
                    __syncwarp();
                    if (++ai == NA) { ai = 0; ++an; }
                }
            }
        } else {
            // ===================== epilogue warps: TMEM -> (registers | int32 tile in shared memory) =====================
            const int ew = warp;
            ev_role = 2;
            if (ew == 0) mark(10);
            const int m = ew * 32 + lane;
            const int et = threadIdx.x;  // 0..127
            const int64_t n = (m < p.R) ? (int64_t) rt * p.R + m : (int64_t) 1 << 40;   // rows beyond the virtual strip: never valid
            const int64_t t0 = (int64_t) tt * NT;
            const int32_t *pre = p.g.act.gpre + (t0 + et) * (p.G + 1);
            float *const out0 = (OUT == 0) ? ten.out : p.g.outs[0];
            if (et < NT) s_sc[et] = __ldg(p.g.act.scales + t0 + et);
            const uint32_t ta0 = tmem_base + ((uint32_t) (ew * 32) << 16);
            if (fast) {
                // the +1 offset of the u8 codes over this CTA's whole slice
                if (et < NT) s_gs[et] = __ldg(pre + g1) - __ldg(pre + g0);
                ptx::bar_sync(1, 128);
                const int nsets = (nblk >= 2 && p.issuers == 2) ? 2 : 1;   // the odd-block issuer had work: its set counts
                WAIT(1, 200, ptx::smem_u32(&s_accfull[0]), 0);
                if (nsets == 2) WAIT(1, 200, ptx::smem_u32(&s_accfull[1]), 0);
                ptx::tc_fence_after();
                if (ew == 0) mark(0);
                const bool ws = (ten.w_scale != 1.0f);
                const int tvalid = (int) min((int64_t) NT, p.g.T - t0);
#pragma unroll 1
                for (int c0 = 0; c0 < NT; c0 += 16) {
                    // sum over the issuers' accumulator sets; MASK: accumulator i holds 4^i x its share.  All loads of a
                    // set are in flight before the one wait (8 dependent load round trips cost ~1000 clk of tail).
                    uint32_t va[16];
#pragma unroll
                    for (int i = 0; i < 16; i++) va[i] = 0;
                    for (int set = 0; set < nsets; set++) {
                        uint32_t v[ACCS][16];
#pragma unroll
                        for (int i = 0; i < ACCS; i++)
                            if (!ABL(4)) ptx::tmem_ld_32x32b_x16(ta0 + (set * ACCS + i) * NT + c0, v[i]);
                        epi::tmem_ld_wait_dep(v[0]);
#pragma unroll
                        for (int i = 1; i < ACCS; i++) epi::reg_fence(v[i]);
#pragma unroll
                        for (int i = 0; i < ACCS; i++)
#pragma unroll
                            for (int j = 0; j < 16; j++) va[j] += (uint32_t) ((int32_t) v[i][j] >> (2 * i));
                    }
                    if (S == 1) {
                        epi::fast_chunk<OUT, false>(va, c0, s_gs, s_sc, ws, ten.w_scale, out0, p.g.outs, p.g.n_out, t0 * ten.ldo + n,
                                                    ten.ldo, tvalid, n < ten.N, nullptr, p.g.vec_ok != 0);
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; i++) red[(c0 + i) * 128 + m] = (int32_t) va[i] - s_gs[c0 + i];
                    }
                }
            } else {
                // exact path: drain at every K-group boundary, int16 wrap, running totals in the shared-memory tile
                for (int g = g0; g < g1; g++) {
                    const uint32_t gi = (uint32_t) (g - g0), ab = gi & 1;
                    if (et < NT) s_gs[et] = __ldg(pre + g + 1) - __ldg(pre + g);
                    ptx::bar_sync(1, 128);
                    WAIT(1, 100, ptx::smem_u32(&s_accfull[ab]), (gi >> 1) & 1);
                    ptx::tc_fence_after();
#pragma unroll 1
                    for (int c0 = 0; c0 < NT; c0 += 16) {
                        int32_t acc[16];
#pragma unroll
                        for (int i = 0; i < 16; i++) acc[i] = 0;
#pragma unroll
                        for (int i = 0; i < ACCS; i++) {
                            uint32_t v[16];
                            ptx::tmem_ld_32x32b_x16(ta0 + (ab * ACCS + i) * NT + c0, v);
                            epi::tmem_ld_wait_dep(v);
#pragma unroll
                            for (int j = 0; j < 16; j++) acc[j] += (int32_t) v[j] >> (2 * i);
                        }
#pragma unroll
                        for (int i = 0; i < 16; i++) {
                            const int32_t x = wrap16(acc[i] - s_gs[c0 + i]);
                            int32_t *r = red + (c0 + i) * 128 + m;
                            *r = (gi == 0) ? x : *r + x;
                        }
                    }
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_accempty[ab]));
                    ptx::bar_sync(1, 128);  // s_gs is rewritten for the next group
                }
                if (S == 1 && n < ten.N) {
#pragma unroll 1
                    for (int c = 0; c < NT; c++) {
                        if (t0 + c < p.g.T) {
                            float v = (float) red[c * 128 + m] * s_sc[c];
                            if (ten.w_scale != 1.0f) v *= ten.w_scale;
                            store_out<OUT>(out0, p.g.outs, p.g.n_out, (t0 + c) * ten.ldo + n, v);
                        }
                    }
                }
            }
        }
    }

    if (warp == 0) { ev_role = 2; mark(1); }
    if (S > 1) {
        // ---- split strip: the S partial tiles meet in distributed shared memory; CTA r finishes token slice r ----
        ptx::cluster_sync();
        if (warp < 4) {
            const int m = warp * 32 + lane;
            const int64_t n = (m < p.R) ? (int64_t) rt * p.R + m : (int64_t) 1 << 40;   // rows beyond the virtual strip: never valid
            const int64_t t0 = (int64_t) tt * NT;
            float *const out0 = (OUT == 0) ? ten.out : p.g.outs[0];
            const int c_lo = (rank * NT) / S, c_hi = ((rank + 1) * NT) / S;
            uint32_t peer[8];
#pragma unroll
            for (int r = 0; r < 8; r++) peer[r] = ptx::mapa(red_base + m * 4, (uint32_t) min(r, S - 1));
#pragma unroll 1
            for (int c = c_lo; c < c_hi; c++) {
                int32_t v[8];
#pragma unroll
                for (int r = 0; r < 8; r++) v[r] = (r < S) ? ptx::ld_shared_cluster_s32(peer[r] + c * 512) : 0;
                const int32_t sum = ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
                if (n < ten.N && t0 + c < p.g.T) {
                    float f = (float) sum * s_sc[c];
                    if (ten.w_scale != 1.0f) f *= ten.w_scale;
                    store_out<OUT>(out0, p.g.outs, p.g.n_out, (t0 + c) * ten.ldo + n, f);
                }
            }
        }
        ptx::cluster_sync();   // nobody's shared memory goes away while a peer still reads it
    }

    // ---- teardown ----
    if (warp == 0) mark(2);
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == W_ACT) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(s_tmem_base, TMEM_COLS);
    }
}

constexpr int SMEM_BUDGET = 216 * 1024;

template <int NT, int DT>
static size_t smem_bytes(int nas, int asb) {
    return (size_t) nas * asb * NT * 128 + (size_t) DT * 4 * DEC_THREADS * 16 + (size_t) NT * 512 + 1024;
}

template <int NT, int OUT, int DT, bool DBG = false>
static cudaError_t launch_out(Params &p, cudaStream_t s) {
    // activation ring: stages of >= 16 KB and >= 4 sub-blocks (two A blocks: one per MMA issuer), as many as the slice
    // needs / the budget allows
    p.ASB = 16384 / (NT * 128);
    if (p.ASB < 4) p.ASB = 4;
    p.asb_shift = p.ASB == 8 ? 3 : 2;
    const int max_groups = (p.G + p.S - 1) / p.S;
    const int need = (max_groups * p.gsub + p.ASB - 1) / p.ASB;
    int nas = need < MAX_AS ? need : MAX_AS;
    if (nas < 1) nas = 1;
    while (nas > 2 && (int) smem_bytes<NT, DT>(nas, p.ASB) > SMEM_BUDGET) --nas;
    const int forced = knobs().tcs_nas;
    if (forced > 0 && forced <= MAX_AS) nas = forced;
    p.NAS = nas;
    const size_t smem = smem_bytes<NT, DT>(p.NAS, p.ASB);
    auto kern = kernel<NT, OUT, DT, DBG>;
    static int smem_set[MAX_DEVICES] = {};
    cudaError_t e = ensure_dyn_smem(kern, smem, p.g.device, smem_set);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) (p.RT * p.S), (unsigned) (p.g.act.Tpad / NT));
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem;
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
    return cudaLaunchKernelEx(&cfg, kern, p);
}

template <int NT, int DT>
static cudaError_t launch_one(Params &p, cudaStream_t s) {
    if (p.g.n_out == 0) {
        if (DT == 3 && NT <= 32 && p.dbg) return launch_out<NT, 0, 3, true>(p, s);   // timeline build (VLUT_TC_DEBUG=1)
        return launch_out<NT, 0, DT>(p, s);
    }
    return p.g.multicast ? launch_out<NT, 2, DT>(p, s) : launch_out<NT, 1, DT>(p, s);
}

// CTAs per strip (S) and rows per strip (R).  Whole K-groups per CTA, at most 8 (portable cluster size), one CTA per SM.
// Measured grid (scripts/grid_tcs.sh, profiles/r2_tcs_split_rows.txt), in units of one 512-k K-group of streaming
// (~0.30 us): a launch costs ~6 units whatever it does, a split strip another ~8 + 0.7 S (cluster launch, cluster
// barrier, distributed-shared-memory reduction) -- so S > 1 only pays for long K (14336 -> 4096: S = 4, 10.3 -> 7.2 us);
// round 1's model charged 0.6 units for a split and picked S = 2..4 for 4096 -> 4096 (5.3 us; S = 1: 4.2 us), 3200 -> 8640
// (6.5 us; 4.0 us) and 2048 -> 8192 (5.9 us; 3.1 us).  R < 128 ("virtual strips": rows [rt R, rt R + R) of the tensor, so
// that e.g. 14336 rows give 144 CTAs on 148 SMs instead of 112) measured 0.3-0.5 us SLOWER at every shape although each
// CTA streams less: kept behind VLUT_TCS_ROWS, not selected.
// Sets p.S, p.R, p.RT and the first (virtual) strip of every tensor of the batch.
static void pick_split_rows(Params &p, int token_tiles) {
    const Knobs &kn = knobs();
    const int smax = p.G < 8 ? p.G : 8;
    const int slots = p.g.sm_count;
    auto strips_of = [&](int R) {
        int64_t n = 0;
        for (int i = 0; i < p.g.n_tens; i++) n += ceil_div(p.g.tens[i].N, R);
        return n;
    };
    int bestS = 1, bestR = TILE_ROWS;
    double best_cost = 1e30;
    for (int s = 1; s <= smax; s++) {
        if (kn.tcs_split > 0 && s != (kn.tcs_split < smax ? kn.tcs_split : smax)) continue;
        if (kn.tcs_split <= 0 && (s & (s - 1)) != 0) continue;   // clusters of 1, 2, 4, 8 (S = 6 on 24 strips measured 2x slower than S = 8 on 16)
        for (int R = TILE_ROWS; R >= 64; R -= 4) {
            if (kn.tcs_rows > 0 ? R != kn.tcs_rows : R != TILE_ROWS) continue;
            const int64_t ctas = strips_of(R) * s * token_tiles;
            if (ctas > slots && s > 1 && kn.tcs_split <= 0) continue;
            const int max_groups = (p.G + s - 1) / s;
            const int waves = (int) ((ctas + slots - 1) / slots);
            // units of one 512-k group of streaming (~0.30 us); an I1 group is 640 k
            const double cost = (double) max_groups * (p.gsub / 4.0) * waves * ((double) R / TILE_ROWS) + 6.0 +
                                (s > 1 ? 8.0 + 0.7 * s : 0.0);
            if (cost < best_cost - 1e-9) { best_cost = cost; bestS = s; bestR = R; }
        }
    }
    p.S = bestS;
    p.R = bestR;
    int strips = 0;
    for (int i = 0; i < p.g.n_tens; i++) {
        p.g.tens[i].strip0 = strips;
        strips += (int) ceil_div(p.g.tens[i].N, p.R);
    }
    p.g.tens[p.g.n_tens].strip0 = strips;   // sentinel
    p.RT = strips;
}

}  // namespace tcs

bool gemm_tcs_supported(int type, int64_t K, int64_t N, int64_t T) {
    return type_valid(type) && T >= 1 && T <= 64 && K >= 4 && N >= 1 && ceil_div(K, group_size(type)) <= 64;
}

cudaError_t launch_gemm_tcs(const GemmArgs &g, cudaStream_t s, int *launches) {
    if (g.act.sw_tile != g.act.NT) return cudaErrorInvalidValue;  // wants full-height swizzled activation tiles
    const int NT = g.act.NT;
    if (NT > 64) return cudaErrorInvalidValue;
    tcs::Params p{};
    p.g = g;
    p.KB = (int) (g.act.Kpad / 128);
    p.gsub = g.act.gsize / 128;
    p.G = g.act.G;
    const Knobs &kn = knobs();
    p.allow_fast = kn.tc_fast;
    p.pdl = kn.pdl;
    tcs::pick_split_rows(p, (int) (g.act.Tpad / NT));   // p.S, p.R, p.RT, virtual strips of the batch
    p.issuers = kn.tcs_issuers;
    p.pre = kn.tcs_pre >= 0 ? kn.tcs_pre : 8;   // default: the whole ring up front (staggering measured slower: 6.0 vs 5.5 us)
    p.dbg = kn.tc_debug ? debug_buffer(s, 0) : nullptr;
    p.dbg_cta = kn.tcs_dbg_cta;
    p.dbg_mask = kn.tcs_dbg_mask;
    if (launches) *launches += 1;
    // ring slots per decode thread: 4 phases x DT blocks = DT x 32 KB of packed weights in flight per CTA
    const int deep = kn.tcs_depth;
    switch (NT) {
        case 16: return deep >= 4 ? tcs::launch_one<16, 4>(p, s) : deep <= 2 ? tcs::launch_one<16, 2>(p, s) : tcs::launch_one<16, 3>(p, s);
        case 32: return deep <= 2 ? tcs::launch_one<32, 2>(p, s) : tcs::launch_one<32, 3>(p, s);
        case 64: return deep <= 2 ? tcs::launch_one<64, 2>(p, s) : tcs::launch_one<64, 3>(p, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace vlut
