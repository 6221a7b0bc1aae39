// gemm_tcs.cu -- VLUT_ALGO_TCS: the small-token-count Vec-LUT GEMM (T <= 64), built to STREAM the packed weights from HBM.
//
// Same arithmetic and exactness contract as gemm_tc4.cu (tcgen05 kind::i8, u8 weight codes x s8 activations -> s32 in TMEM,
// per-K-group int16 wrap on the exact path, fused fp32 epilogue), reorganised around what the
// timelines of that kernel showed at T <= 64: the HBM stream, not arithmetic, must set the pace, and every hand-off in
// the chain  TMA -> mbarrier -> LDS -> unpack -> tcgen05.st -> MMA  costs a memory or barrier round trip.
//
//   * WEIGHTS never pass through a producer warp or an mbarrier.  Decode thread (q, h, lane) owns weight row
//     m = 32q + lane of the strip and sub-block h of every 256-k A block; it fetches its own 2 x 16 bytes per block with
//     cp.async (LDGSTS, coalesced 512 B per warp instruction, L2 evict_first) into a private ring of D blocks in shared
//     memory and reads them back itself (cp.async.wait_group: no cross-thread synchronisation at all).  D x 8 KB per CTA
//     are in flight from the first instruction of the kernel -- before griddepcontrol.wait, i.e. while the quantise
//     kernel is still running -- which is what saturating HBM needs (Little: 6.5 TB/s x ~1 us over <= 148 CTAs).
//   * K is SPLIT over the CTAs of a thread-block cluster (S = 1..8 CTAs per 128-row strip, whole K-groups each) so that
//     shapes with few strips (4096 -> 4096: 32) still put >= 128 CTAs on the weight stream.  Partial sums meet in
//     distributed shared memory: every CTA parks its int32 tile in its own shared memory, one cluster barrier, then CTA r
//     reduces and stores the token slice r of the tile (ld.shared::cluster).  Integer partials: order-independent, exact.
//   * 256 TMEM columns and <= ~110 KB of shared memory per CTA at NT <= 32: two CTAs share an SM, so a successor kernel
//     (PDL) or the second CTA of a split strip overlaps its prologue and weight prefetch with this one's drain.
//
//   warps 0-3   epilogue (TMEM lane quadrant = warp)        warps 4-11  decode (quadrant q = (w-4)%4, sub-block h = (w-4)/4)
//   warp 12     TMEM allocator + activation producer (TMA)  warp 13     MMA issuer
#include <stdlib.h>

#include "epilogue.cuh"
#include "ptx.cuh"
#include "vlut_common.cuh"

#ifndef VLUT_TCS_DEC_MARKS
#define VLUT_TCS_DEC_MARKS 31
#endif
#define DECM(bit) ((VLUT_TCS_DEC_MARKS & (bit)) != 0)

namespace vlut {

long long *debug_buffer(cudaStream_t s, int skip);  // gemm_tc.cu

namespace tcs {

constexpr int THREADS = 14 * 32;
constexpr int W_DEC0 = 4, NDEC = 8, W_ACT = 12, W_MMA = 13;
constexpr int SPA = 2;               // 128-k sub-blocks per A block (decode -> MMA hand-off): 256 k
constexpr int A_COLS = SPA * 32;     // TMEM columns of one A stage
constexpr int TMEM_COLS = 256;
constexpr int MAX_AS = 8;            // activation stages (ring)
constexpr int DEC_THREADS = NDEC * 32;
constexpr int W_BLOCK_BYTES = 2 * SPA * TILE_BYTES;  // packed bytes of one A block of a strip: 8 KB

struct Params {
    GemmArgs g;
    int KB, G, gsub, RT;
    int S;          // CTAs per strip (cluster size): K-groups split [r*G/S, (r+1)*G/S)
    int NAS, ASB;   // activation ring: NAS stages of ASB 128-k sub-blocks
    int allow_fast, pdl;
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

template <int NT, int OUT, int D, bool DBG>
__global__ void __launch_bounds__(THREADS, (NT <= 32) ? 2 : 1) kernel(const __grid_constant__ Params p) {
    constexpr int NA = (NT == 64) ? 2 : 3;            // A stages in TMEM
    constexpr uint32_t A_COL0 = 2 * NT;               // two accumulators of NT columns, then the A stages
    static_assert(A_COL0 + NA * A_COLS <= TMEM_COLS, "TMEM budget");
    constexpr int ACT_BYTES = NT * 128;               // one 128-k sub-block of the token tile, pre-swizzled by the quantiser
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_actfull[MAX_AS], s_actempty[MAX_AS];
    __shared__ __align__(8) uint64_t s_afull[NA], s_aempty[NA];
    __shared__ __align__(8) uint64_t s_accfull[2], s_accempty[2];
    __shared__ uint32_t s_tmem_base;
    __shared__ __align__(16) int32_t s_gs[NT];
    __shared__ __align__(16) float s_sc[NT];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = p.S;
    const int rank = (S > 1) ? (int) ptx::cluster_ctarank() : 0;
    const int rt = blockIdx.x / S;
    const int tt = blockIdx.y;
    const int g0 = (rank * p.G) / S, g1 = ((rank + 1) * p.G) / S;   // this CTA's K-groups
    const int kb_begin = g0 * p.gsub, kb_end = min(g1 * p.gsub, p.KB);
    const int nsub_slice = kb_end - kb_begin;

    // dynamic shared memory: [activation ring][weight ring: D blocks x 2 chunks x 256 threads x 16 B][int32 tile NT x 128]
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t act_base = smem_base;
    const uint32_t wring = act_base + (uint32_t) (p.NAS * p.ASB) * ACT_BYTES;
    const uint32_t red_base = wring + D * W_BLOCK_BYTES;
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

    // let the next kernel of the stream start its prologue (and its own weight prefetch) as early as resources allow
    if (p.pdl) ptx::pdl_launch_dependents();

    if (warp >= W_DEC0 && warp < W_DEC0 + NDEC) {
        // ===================== decode: HBM -> (cp.async ring) -> registers -> u8 A operand in TMEM =====================
        const int q = (warp - W_DEC0) & 3, h = (warp - W_DEC0) >> 2;
        const int m = q * 32 + lane;
        const int dt = (warp - W_DEC0) * 32 + lane;
        ev_role = 0;
        const bool dmark = (warp == W_DEC0);
        if (dmark && DECM(16)) mark(10);
        const uint64_t pol = ptx::policy_evict_first();
        // A blocks of this CTA: pairs of 128-k sub-blocks over the K slice, whatever the K-groups are (group boundaries only
        // concern the MMA issuer and the epilogue); this thread unpacks sub-block h of every block: 2 x 16 B, 8 KB further
        // on for every block
        const int nblk = (nsub_slice + SPA - 1) / SPA;
        const uint8_t *wsrc = p.g.w + ((int64_t) rt * (2 * p.KB) + 2 * (kb_begin + h)) * TILE_BYTES + m * 16;
        const uint32_t wdst = wring + dt * 16;
        auto have = [&](int b) { return SPA * b + h < nsub_slice; };
        auto issue = [&](int b, int slot) {   // this thread's 32 bytes of block b -> ring slot; always one commit group
            if (b < nblk && have(b)) {
                const uint8_t *src = wsrc + (int64_t) b * W_BLOCK_BYTES;
                ptx::cp_async16_hint(wdst + (slot * 2 + 0) * (DEC_THREADS * 16), src, pol);
                ptx::cp_async16_hint(wdst + (slot * 2 + 1) * (DEC_THREADS * 16), src + TILE_BYTES, pol);
            }
            ptx::cp_async_commit();
        };
        // weights are static data: the stream starts now, whatever the predecessor kernel (quantise) is still doing
#pragma unroll
        for (int i = 0; i < D; i++) issue(i, i);
        // barriers / TMEM address are set up by warp 12 meanwhile
        ptx::bar_sync(2, THREADS);
        ptx::tc_fence_after();
        const uint32_t tmem_base = s_tmem_base;
        const uint32_t taddr0 = tmem_base + ((uint32_t) (q * 32) << 16) + A_COL0 + h * 32;
        int slot = 0, ai = 0;
        uint32_t aph = 0;
#pragma unroll 1
        for (int b = 0; b < nblk; b++) {
            ptx::cp_async_wait<D - 1>();
            if (dmark && DECM(1)) mark(0);
            const bool hv = have(b);
            uint4 cw[2];
            if (hv) {
                cw[0] = ptx::lds128(wdst + (slot * 2 + 0) * (DEC_THREADS * 16));
                cw[1] = ptx::lds128(wdst + (slot * 2 + 1) * (DEC_THREADS * 16));
            }
            issue(b + D, slot);   // refill the slot just read (same thread: program order)
            ptx::mbar_wait(ptx::smem_u32(&s_aempty[ai]), aph ^ 1);
            ptx::tc_fence_after();
            if (dmark && DECM(2)) mark(1);
            if (hv) {
#pragma unroll
                for (int c = 0; c < 2; c++) {
                    const uint32_t wv[4] = {cw[c].x, cw[c].y, cw[c].z, cw[c].w};
                    uint32_t r[16];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        // shifts as multiply-high on the FMA pipe, masks on the ALU pipe
                        r[4 * j + 0] = wv[j] & 0x03030303u;
                        r[4 * j + 1] = __umulhi(wv[j], 1u << 30) & 0x03030303u;
                        r[4 * j + 2] = __umulhi(wv[j], 1u << 28) & 0x03030303u;
                        r[4 * j + 3] = __umulhi(wv[j], 1u << 26) & 0x03030303u;
                    }
                    ptx::tmem_st_32x32b_x16(taddr0 + ai * A_COLS + c * 16, r);
                }
                if (dmark && DECM(4)) mark(2);
                ptx::tmem_st_wait();
            }
            if (dmark && DECM(8)) mark(3);
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_afull[ai]));
            __syncwarp();
            if (++ai == NA) { ai = 0; aph ^= 1; }
            if (++slot == D) slot = 0;
        }
        ptx::cp_async_wait<0>();
    } else {
        if (warp == W_ACT) {
            if (lane == 0) {
                for (int s = 0; s < p.NAS; s++) {
                    ptx::mbar_init(ptx::smem_u32(&s_actfull[s]), 1);
                    ptx::mbar_init(ptx::smem_u32(&s_actempty[s]), 1);
                }
                for (int d = 0; d < NA; d++) {
                    ptx::mbar_init(ptx::smem_u32(&s_afull[d]), NDEC);
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
            const int n_as = (nsub_slice + p.ASB - 1) / p.ASB;
            const int8_t *src0 = p.g.act.q_sw + ((int64_t) tt * p.KB + kb_begin) * ACT_BYTES;
            for (int asi = 0, ring = 0, ph = 0; asi < n_as; asi++) {
                ptx::mbar_wait(ptx::smem_u32(&s_actempty[ring]), ph ^ 1);
                if (ptx::elect_one()) {
                    const uint32_t bytes = (uint32_t) min(p.ASB, nsub_slice - asi * p.ASB) * ACT_BYTES;
                    const uint32_t full = ptx::smem_u32(&s_actfull[ring]);
                    ptx::mbar_arrive_expect_tx(full, bytes);
                    ptx::bulk_g2s_hint(act_base + (uint32_t) (ring * p.ASB) * ACT_BYTES, src0 + (int64_t) asi * p.ASB * ACT_BYTES,
                                       bytes, full, pol);
                }
                __syncwarp();
                mark(1);
                if (++ring == p.NAS) { ring = 0; ph ^= 1; }
            }
        } else if (warp == W_MMA) {
            // ===================== MMA issuer =====================
            constexpr uint32_t idesc = make_idesc(NT);
            constexpr uint32_t SUBD = ACT_BYTES >> 4;   // shared-memory descriptor units per 128-k sub-block
            ev_role = 1;
            mark(10);
            const uint64_t desc0 = make_smem_desc(act_base);
            const int nblk = (nsub_slice + SPA - 1) / SPA;
            int ai = 0, ring = 0, asr = 0;   // A stage; activation ring stage and sub-block position inside it
            uint32_t aph = 0, as_ph = 0;
            if (fast) {
                // Hot loop.  One thread feeds the tensor pipe and its scalar instruction stream IS the critical path at
                // N = 16 (scripts/ubench/umma_contend.cu: 40-65 clk per MMA with per-MMA operand set-up and a re-election
                // per sub-block, against ~18 clk for the MMA itself): one wait, one election and one straight-line burst
                // of 8 MMAs + commits per A block, every operand an increment of a uniform register.
                const uint32_t d_tmem = tmem_base;
#pragma unroll 1
                for (int b = 0; b < nblk; b++) {
                    ptx::mbar_wait(ptx::smem_u32(&s_afull[ai]), aph);
                    if (asr == 0) ptx::mbar_wait(ptx::smem_u32(&s_actfull[ring]), as_ph);
                    ptx::tc_fence_after();
                    mark(0);
                    const bool two = SPA * b + 1 < nsub_slice;
                    const bool last = (b == nblk - 1);
                    const bool rel_as = (asr + SPA >= p.ASB) || last;
                    if (ptx::elect_one()) {
                        const uint32_t a_tmem = tmem_base + A_COL0 + ai * A_COLS;
                        const uint64_t bd = desc0 + (uint64_t) ((uint32_t) (ring * p.ASB + asr) * SUBD);
#pragma unroll
                        for (int ks = 0; ks < 4; ks++)
                            ptx::umma_i8_ts(d_tmem, a_tmem + ks * 8, bd + (uint64_t) (ks * 2), idesc, (b > 0 || ks > 0) ? 1u : 0u);
                        if (two) {
#pragma unroll
                            for (int ks = 0; ks < 4; ks++)
                                ptx::umma_i8_ts(d_tmem, a_tmem + 32 + ks * 8, bd + (uint64_t) (SUBD + ks * 2), idesc, 1u);
                        }
                        ptx::umma_commit(ptx::smem_u32(&s_aempty[ai]));
                        if (rel_as) ptx::umma_commit(ptx::smem_u32(&s_actempty[ring]));
                        if (last) ptx::umma_commit(ptx::smem_u32(&s_accfull[0]));
                    }
                    __syncwarp();
                    mark(1);
                    if (++ai == NA) { ai = 0; aph ^= 1; }
                    asr += SPA;
                    if (rel_as) {
                        asr = 0;
                        if (++ring == p.NAS) { ring = 0; as_ph ^= 1; }
                    }
                }
            } else {
                // Exact path (some token of the tile can overflow int16 inside a K-group): one accumulator drain per K-group,
                // the two accumulators alternate.  Generic per-sub-block loop; group boundaries may fall inside an A block.
                uint32_t ndrain = 0, ab = 0;
                int g = g0, gpos = 0;          // current K-group and sub-block position inside it
                bool fresh = true;
#pragma unroll 1
                for (int b = 0; b < nblk; b++) {
                    ptx::mbar_wait(ptx::smem_u32(&s_afull[ai]), aph);
                    ptx::tc_fence_after();
                    const int ns = min(SPA, nsub_slice - SPA * b);
                    for (int s = 0; s < ns; s++) {
                        const int kb = kb_begin + SPA * b + s;
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
                            const uint32_t d_tmem = tmem_base + ab * NT;
                            const uint32_t a_tmem = tmem_base + A_COL0 + ai * A_COLS + s * 32;
                            const uint64_t bd = desc0 + (uint64_t) ((uint32_t) (ring * p.ASB + asr) * SUBD);
#pragma unroll
                            for (int ks = 0; ks < 4; ks++)
                                ptx::umma_i8_ts(d_tmem, a_tmem + ks * 8, bd + (uint64_t) (ks * 2), idesc, (fresh && ks == 0) ? 0u : 1u);
                            if (rel_as) ptx::umma_commit(ptx::smem_u32(&s_actempty[ring]));
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
                    if (++ai == NA) { ai = 0; aph ^= 1; }
                }
            }
        } else {
            // ===================== epilogue warps: TMEM -> (registers | int32 tile in shared memory) =====================
            const int ew = warp;
            ev_role = 2;
            if (ew == 0) mark(10);
            const int m = ew * 32 + lane;
            const int et = threadIdx.x;  // 0..127
            const int64_t n = (int64_t) rt * TILE_ROWS + m;
            const int64_t t0 = (int64_t) tt * NT;
            const int32_t *pre = p.g.act.gpre + (t0 + et) * (p.G + 1);
            float *const out0 = (OUT == 0) ? p.g.out : p.g.outs[0];
            if (et < NT) s_sc[et] = __ldg(p.g.act.scales + t0 + et);
            const uint32_t ta0 = tmem_base + ((uint32_t) (ew * 32) << 16);
            if (fast) {
                // the +1 offset of the u8 codes over this CTA's whole slice
                if (et < NT) s_gs[et] = __ldg(pre + g1) - __ldg(pre + g0);
                ptx::bar_sync(1, 128);
                ptx::mbar_wait(ptx::smem_u32(&s_accfull[0]), 0);
                ptx::tc_fence_after();
                if (ew == 0) mark(0);
                if (S == 1) {
                    epi::drain_fast<NT, OUT, false>(ta0, s_gs, s_sc, p.g.w_scale, out0, p.g.outs, p.g.n_out, t0 * p.g.ldo + n, p.g.ldo,
                                                    (int) min((int64_t) NT, p.g.T - t0), n < p.g.N, nullptr);
                } else {
#pragma unroll 1
                    for (int c0 = 0; c0 < NT; c0 += 16) {
                        uint32_t v[16];
                        ptx::tmem_ld_32x32b_x16(ta0 + c0, v);
                        epi::tmem_ld_wait_dep(v);
#pragma unroll
                        for (int i = 0; i < 16; i++) red[(c0 + i) * 128 + m] = (int32_t) v[i] - s_gs[c0 + i];
                    }
                }
            } else {
                // exact path: drain at every K-group boundary, int16 wrap, running totals in the shared-memory tile
                for (int g = g0; g < g1; g++) {
                    const uint32_t gi = (uint32_t) (g - g0), ab = gi & 1;
                    if (et < NT) s_gs[et] = __ldg(pre + g + 1) - __ldg(pre + g);
                    ptx::bar_sync(1, 128);
                    ptx::mbar_wait(ptx::smem_u32(&s_accfull[ab]), (gi >> 1) & 1);
                    ptx::tc_fence_after();
#pragma unroll 1
                    for (int c0 = 0; c0 < NT; c0 += 16) {
                        uint32_t v[16];
                        ptx::tmem_ld_32x32b_x16(ta0 + ab * NT + c0, v);
                        epi::tmem_ld_wait_dep(v);
#pragma unroll
                        for (int i = 0; i < 16; i++) {
                            const int32_t x = wrap16((int32_t) v[i] - s_gs[c0 + i]);
                            int32_t *r = red + (c0 + i) * 128 + m;
                            *r = (gi == 0) ? x : *r + x;
                        }
                    }
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&s_accempty[ab]));
                    ptx::bar_sync(1, 128);  // s_gs is rewritten for the next group
                }
                if (S == 1 && n < p.g.N) {
#pragma unroll 1
                    for (int c = 0; c < NT; c++) {
                        if (t0 + c < p.g.T) {
                            float v = (float) red[c * 128 + m] * s_sc[c];
                            if (p.g.w_scale != 1.0f) v *= p.g.w_scale;
                            store_out<OUT>(out0, p.g.outs, p.g.n_out, (t0 + c) * p.g.ldo + n, v);
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
            const int64_t n = (int64_t) rt * TILE_ROWS + m;
            const int64_t t0 = (int64_t) tt * NT;
            float *const out0 = (OUT == 0) ? p.g.out : p.g.outs[0];
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
                if (n < p.g.N && t0 + c < p.g.T) {
                    float f = (float) sum * s_sc[c];
                    if (p.g.w_scale != 1.0f) f *= p.g.w_scale;
                    store_out<OUT>(out0, p.g.outs, p.g.n_out, (t0 + c) * p.g.ldo + n, f);
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

static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return v ? atoi(v) : dflt;
}

constexpr int SMEM_PAIR_BUDGET = 112 * 1024;   // two CTAs per SM
constexpr int SMEM_SOLO_BUDGET = 208 * 1024;

template <int NT, int D>
static size_t smem_bytes(int nas, int asb) {
    return (size_t) nas * asb * NT * 128 + (size_t) D * W_BLOCK_BYTES + (size_t) NT * 512 + 1024;
}

template <int NT, int OUT, int D, bool DBG = false>
static cudaError_t launch_out(Params &p, cudaStream_t s) {
    // activation ring: stages of ~16 KB, as many as the slice needs / the budget allows
    p.ASB = 16384 / (NT * 128);
    const int max_groups = (p.G + p.S - 1) / p.S;
    const int need = (max_groups * p.gsub + p.ASB - 1) / p.ASB;
    const bool pair_ok = (NT <= 32);
    const bool want_pair = pair_ok && ((int64_t) p.RT * p.S > p.g.sm_count || env_int("VLUT_TCS_PAIR", 1));
    int nas = need < MAX_AS ? need : MAX_AS;
    if (nas < 2) nas = need < 2 ? need : 2;
    if (nas < 1) nas = 1;
    const int budget = want_pair ? SMEM_PAIR_BUDGET : SMEM_SOLO_BUDGET;
    while (nas > 2 && (int) smem_bytes<NT, D>(nas, p.ASB) > budget) --nas;
    const int forced = env_int("VLUT_TCS_NAS", 0);
    if (forced > 0 && forced <= MAX_AS) nas = forced;
    p.NAS = nas;
    const size_t smem = smem_bytes<NT, D>(p.NAS, p.ASB);
    auto kern = kernel<NT, OUT, D, DBG>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
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

template <int NT, int D>
static cudaError_t launch_one(Params &p, cudaStream_t s) {
    if (p.g.n_out == 0) {
        if (D == 8 && NT <= 32 && p.dbg) return launch_out<NT, 0, 8, true>(p, s);   // timeline build (VLUT_TC_DEBUG=1)
        return launch_out<NT, 0, D>(p, s);
    }
    return p.g.multicast ? launch_out<NT, 2, D>(p, s) : launch_out<NT, 1, D>(p, s);
}

// CTAs per strip.  More CTAs = more weight bytes in flight and shorter per-CTA streams; a split strip pays one cluster
// barrier and a distributed-shared-memory reduction.  Whole K-groups per CTA, at most 8 (portable cluster size).
static int pick_split(const Params &p, int NT) {
    const int forced = env_int("VLUT_TCS_SPLIT", 0);
    const int smax = p.G < 8 ? p.G : 8;
    if (forced > 0) return forced < smax ? forced : smax;
    const int slots = p.g.sm_count * (NT <= 32 ? 2 : 1);   // co-resident CTAs
    int best = 1;
    double best_cost = 1e30;
    for (int s = 1; s <= smax; s++) {
        const int64_t ctas = (int64_t) p.RT * s;
        if (ctas > slots && s > 1) break;
        const int max_groups = (p.G + s - 1) / s;
        const int waves = (int) ((ctas + slots - 1) / slots);
        // time ~ the longest per-CTA stream, but never less than the whole matrix at full HBM rate, + a fixed cost per split
        const double stream = (double) max_groups * waves;
        const double floor_ = (double) p.RT * p.G / 120.0;   // ~120 CTA-streams saturate HBM
        const double cost = (stream > floor_ ? stream : floor_) + 0.35 * (s > 1) + 0.05 * s;
        if (cost < best_cost - 1e-9) { best_cost = cost; best = s; }
    }
    return best;
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
    p.RT = (int) (g.Npad / TILE_ROWS);
    p.allow_fast = tcs::env_int("VLUT_TC_FAST", 1);
    p.pdl = tcs::env_int("VLUT_PDL", 1);
    p.S = tcs::pick_split(p, NT);
    p.dbg = tcs::env_int("VLUT_TC_DEBUG", 0) ? debug_buffer(s, 0) : nullptr;
    p.dbg_cta = tcs::env_int("VLUT_TCS_DBG_CTA", 0);
    p.dbg_mask = tcs::env_int("VLUT_TCS_DBG_MASK", 15);
    if (launches) *launches += 1;
    const int deep = tcs::env_int("VLUT_TCS_DEPTH", 8);
    switch (NT) {
        case 16: return deep >= 16 ? tcs::launch_one<16, 16>(p, s) : tcs::launch_one<16, 8>(p, s);
        case 32: return deep >= 16 ? tcs::launch_one<32, 16>(p, s) : tcs::launch_one<32, 8>(p, s);
        case 64: return tcs::launch_one<64, 8>(p, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace vlut
