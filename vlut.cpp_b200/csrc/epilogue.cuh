// epilogue.cuh -- accumulator drain shared by the tcgen05 kernels (gemm_tc4.cu, gemm_tc5.cu).
//
// The first version of the drain spent ~350 SASS instructions per 16 accumulator columns (per-value branches for the
// exact path, per-store token predicates, 64-bit index arithmetic) and was issue-latency bound: one epilogue warp per
// scheduler needed ~18,000 clk per 128-column tile, more than twice the tile's tensor-pipe time
// (profiles/r1_ncu_tc5_T2048_source.txt).  The hot cases -- fast path (one drain per tile run), complete tile -> fp32
// output, or split tile -> int32 partial slot -- are specialised here: branch-free per value, vector shared-memory reads
// of the per-token corrections and scales, the next tcgen05.ld in flight while the current 16 columns are converted and
// stored.  The exact (per-K-group int16 wrap) path keeps the generic code in the kernels.
#pragma once
#include "ptx.cuh"
#include "vlut_common.cuh"

namespace vlut {
namespace epi {

// tcgen05.wait::ld that also "produces" the loaded registers, so no consumer can be scheduled above it
__device__ __forceinline__ void tmem_ld_wait_dep(uint32_t (&v)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                   "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])::"memory");
}

// no instruction: registers loaded by an earlier tcgen05.ld become "defined" here, i.e. after the wait that precedes this
// statement (volatile asm statements keep their order), so no consumer can be scheduled above that wait
__device__ __forceinline__ void reg_fence(uint32_t (&v)[16]) {
    asm volatile(""
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                   "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15])::"memory");
}

// 4 x 4 transpose inside every group of four lanes: on entry lane 4g + j holds a[tt] = (row 4g + j, token tt); on exit it
// holds a[rr] = (row 4g + rr, token j).  Two butterfly stages, 4 shuffles.
__device__ __forceinline__ void transpose4(float (&a)[4], int lane) {
    const bool b0 = lane & 1, b1 = lane & 2;
    {
        const float s0 = b0 ? a[0] : a[1], s1 = b0 ? a[2] : a[3];
        const float r0 = __shfl_xor_sync(0xffffffffu, s0, 1), r1 = __shfl_xor_sync(0xffffffffu, s1, 1);
        if (b0) { a[0] = r0; a[2] = r1; } else { a[1] = r0; a[3] = r1; }
    }
    {
        const float s0 = b1 ? a[0] : a[2], s1 = b1 ? a[1] : a[3];
        const float r0 = __shfl_xor_sync(0xffffffffu, s0, 2), r1 = __shfl_xor_sync(0xffffffffu, s1, 2);
        if (b1) { a[0] = r0; a[1] = r1; } else { a[2] = r0; a[3] = r1; }
    }
}

// Staged scatter (OUT == 3, the CTA-pair kernel's fused all-gather over peer pointers).  Direct epilogue stores to peer
// memory fill NVLink (688 GB/s with this very access pattern, scripts/ubench/peer_store.cu) -- but inside the GEMM they
// took 91 us for a 33 us GEMM + 43 us of link time: the stalled remote stores sit in the SM's load/store path, in front of
// the decode warps' shared-memory loads and the next accumulator's drain.  So the four epilogue warps park 16 tokens x 128
// rows of fp32 in a shared-memory buffer (a ring of SCATTER_NBUF), and one bulk async copy per (token, destination) -- 512 contiguous
// bytes, issued by up to 128 different threads -- carries them out through the TMA engine: the SM's store path stays
// empty and the copies drain under the following MMAs.
// Protocol per 16-token chunk k (buffer k % 2), all 128 epilogue threads:
//   everybody writes buffer k % NBUF (free since the barrier of chunk k - 1) -> fence.proxy.async -> issuers wait until
//   their bulk copies of chunk k - (NBUF - 1) have READ shared memory (so buffer (k + 1) % NBUF is free for the next chunk)
//   -> barrier -> issuers launch the copies of chunk k and commit.  One named barrier per chunk.  With two buffers only
//   8 KB per destination and CTA were in flight: 1.2 MB chip-wide against ~2 MB of NVLink bandwidth x latency.
struct Scatter {
    uint32_t stage;      // shared-memory address of the SCATTER_NBUF 16 x 128 x 4 B buffers
    uint32_t k;          // chunks staged so far
    int et;              // epilogue thread 0..127 == row of the strip
    int row_bytes;       // bytes of the strip's rows that exist (ragged last strip), multiple of 16
    // OUT == 4 (local output through a tensor map): one 2-D TMA store per 16-token chunk, issued by thread 0; the hardware
    // clips rows / tokens beyond the matrix
    const void *tmap;    // CUtensorMap of the output [T][ldo] (box 128 rows x 16 tokens), in kernel parameter space
    int n0, tok;         // first row of the strip; token of the next chunk (set per tile, advanced per chunk)
    __device__ __forceinline__ void init(uint32_t stage_smem, int et_) {
        stage = stage_smem; k = 0; et = et_; row_bytes = 512; tmap = nullptr; n0 = 0; tok = 0;
    }
    __device__ __forceinline__ void finish() const {   // every copy of this thread has been written out
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
};
constexpr int SCATTER_NBUF = 3;   // staging buffers: NBUF - 1 chunks (8 KB per destination each) in flight towards the peers per CTA
constexpr int SCATTER_BYTES = SCATTER_NBUF * 16 * 128 * 4;

// 16 final int32 sums of one weight row (tokens t .. t+15) -> fp32 outputs: out[(t+i) * ldo + n] = (float) x * s[t+i]
// (* w_scale as a second rounding when `ws`).  nv = number of valid tokens among the 16 (>= 16: all); nvalid = this
// lane's row exists.  WARP-COLLECTIVE for the scatter flavours (OUT != 0): every lane of the warp must call it.
//
// Scatter flavours (multi-GPU fused all-gather): the values cross NVLink, where a 4-byte store per lane is a 4-byte
// transaction (7.3 M of them per rank per step at T = 512 filled 31-66 % of the link, round 1).  Lanes are transposed in
// groups of four so that every lane stores FOUR consecutive rows of one token as one 16-byte multimem.st / peer store:
// a warp instruction then writes 4 tokens x 128 contiguous bytes.  Needs 16-byte aligned rows (vec_ok); ragged row tails and
// unaligned destinations take the scalar stores.
template <int OUT>
__device__ __forceinline__ void store16(const int32_t (&x)[16], const float *sc_smem, bool ws, float w_scale, float *out0,
                                        float *const *outs, int n_out, int64_t idx0, int64_t ldo, int nv, bool nvalid,
                                        bool vec_ok = false, Scatter *sc = nullptr) {
    float f[16];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const float4 s = *reinterpret_cast<const float4 *>(sc_smem + 4 * j);
        f[4 * j + 0] = (float) x[4 * j + 0] * s.x;
        f[4 * j + 1] = (float) x[4 * j + 1] * s.y;
        f[4 * j + 2] = (float) x[4 * j + 2] * s.z;
        f[4 * j + 3] = (float) x[4 * j + 3] * s.w;
    }
    if (ws) {
#pragma unroll
        for (int i = 0; i < 16; i++) f[i] *= w_scale;
    }
    constexpr int SO = (OUT == 4) ? 0 : OUT;   // the direct stores of the staged flavours' slow paths
    if (OUT == 4 && sc != nullptr) {
        // local output, staged: 16 st.shared per thread (immediate offsets, no address arithmetic), one named barrier and ONE
        // tensor store per chunk.  (The direct drain spends ~115 instructions per thread and chunk, 48 of them address adds and
        // STG; sixteen 512-byte bulk copies per chunk instead of the tensor store measured slower than the direct stores.)
        const uint32_t buf = sc->stage + (sc->k % SCATTER_NBUF) * (16 * 512);
#pragma unroll
        for (int i = 0; i < 16; i++)
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(buf + i * 512 + sc->et * 4), "f"(f[i]) : "memory");
        ptx::fence_proxy_async();
        if (sc->et == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(SCATTER_NBUF - 2) : "memory");
        ptx::bar_sync(1, 128);
        if (sc->et == 0) {
            if (nv > 0) ptx::tma_store_2d(sc->tmap, sc->n0, sc->tok, buf);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        sc->tok += 16;
        sc->k++;
        return;
    }
    if (OUT == 3 && sc != nullptr) {
        // all 128 epilogue threads, same chunk: idx0 = (token of column 0) * ldo + (first row of the strip + et)
        const int n_dst = n_out;
        // 16 * n_dst copies per chunk, issued by threads spread over all four warps: bulk copies issued by one warp
        // serialise (2 destinations from the 32 lanes of warp 0 moved 8 KB per ~1300 clk)
        const int ncopies = 16 * n_dst, stride = ncopies >= 128 ? 1 : 128 / ncopies;
        const int cidx = sc->et / stride;
        const bool issuer = (sc->et % stride == 0) && cidx < ncopies;
        const uint32_t buf = sc->stage + (sc->k % SCATTER_NBUF) * (16 * 512);
#pragma unroll
        for (int i = 0; i < 16; i++)
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(buf + i * 512 + sc->et * 4), "f"(f[i]) : "memory");
        ptx::fence_proxy_async();
        // chunk k - (NBUF - 1) has left its buffer, which the NEXT chunk will write
        if (issuer) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(SCATTER_NBUF - 2) : "memory");
        ptx::bar_sync(1, 128);
        if (issuer) {
            const int i = cidx & 15, d = cidx >> 4;
            if (i < nv) {
                float *dst = outs[d] + (idx0 - sc->et) + (int64_t) i * ldo;
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(buf + i * 512),
                             "r"(sc->row_bytes)
                             : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        sc->k++;
        return;
    }
    if (vec_ok) {   // (local output only with VLUT_TC_VEC_OUT=1: measured slower there than the 4-byte stores)
        const int lane = threadIdx.x & 31, j = lane & 3;
        // which rows of this lane's group of four exist (ragged tail of the matrix)
        const unsigned gm = (__ballot_sync(0xffffffffu, nvalid) >> (lane & ~3)) & 0xfu;
        // idx of (token c0, row of lane 4g): this lane's idx0 minus its position in the group
        const int64_t idxg = idx0 - j;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            float a[4] = {f[4 * i + 0], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]};
            transpose4(a, lane);
            // now: rows 4g .. 4g+3 of token 4i + j
            const int tok = 4 * i + j;
            if (tok < nv) {
                if (gm == 0xfu) {
                    store_out4<SO>(out0, outs, n_out, idxg + tok * ldo, make_float4(a[0], a[1], a[2], a[3]));
                } else {
#pragma unroll
                    for (int r = 0; r < 4; r++)
                        if ((gm >> r) & 1u) store_out<SO>(out0, outs, n_out, idxg + tok * ldo + r, a[r]);
                }
            }
        }
        return;
    }
    if (!nvalid) return;
    if (nv >= 16) {
        if (SO == 0) {
            // running pointer: one 64-bit add per store instead of a 64-bit multiply-add chain per store
            float *pr = out0 + idx0;
#pragma unroll
            for (int i = 0; i < 16; i++) {
                *pr = f[i];
                pr += ldo;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; i++) store_out<SO>(out0, outs, n_out, idx0 + i * ldo, f[i]);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 16; i++)
            if (i < nv) store_out<SO>(out0, outs, n_out, idx0 + i * ldo, f[i]);
    }
}

// One 16-column chunk of a fast-path drain: subtract the u8-code offset correction, then store (fp32 output or partial slot).
template <int OUT, int TO_SLOT>  // TO_SLOT: 0 fp32 output, 1 store the int32 partial, 2 red.add it into an accumulation tile
__device__ __forceinline__ void fast_chunk(const uint32_t (&v)[16], int c0, const int32_t *s_gs, const float *s_sc, bool ws,
                                           float w_scale, float *out0, float *const *outs, int n_out, int64_t idx0,
                                           int64_t ldo, int tvalid, bool nvalid, int32_t *slot_m, bool vec_ok = false,
                                           Scatter *sc = nullptr) {
    int32_t x[16];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int4 g = *reinterpret_cast<const int4 *>(s_gs + c0 + 4 * j);
        x[4 * j + 0] = (int32_t) v[4 * j + 0] - g.x;
        x[4 * j + 1] = (int32_t) v[4 * j + 1] - g.y;
        x[4 * j + 2] = (int32_t) v[4 * j + 2] - g.z;
        x[4 * j + 3] = (int32_t) v[4 * j + 3] - g.w;
    }
    if (TO_SLOT == 2) {
#pragma unroll
        for (int i = 0; i < 16; i++) atomicAdd(slot_m + (c0 + i) * 128, x[i]);  // result unused: RED, resolved in L2
    } else if (TO_SLOT == 1) {
#pragma unroll
        for (int i = 0; i < 16; i++) __stcg(slot_m + (c0 + i) * 128, x[i]);
    } else {
        store16<OUT>(x, s_sc + c0, ws, w_scale, out0, outs, n_out, idx0 + (int64_t) c0 * ldo, ldo, tvalid - c0, nvalid, vec_ok, sc);
    }
}

// Fast-path drain of one accumulator (NT columns of this thread's TMEM lane).  `ta` = TMEM address of column 0 for this
// warp's lane quadrant; s_gs / s_sc = per-token correction and scale (shared memory, 16-byte aligned); idx0 = t0 * ldo + n;
// tvalid = T - t0; slot_m = partial slot + m (TO_SLOT).  Warp-collective: every lane must call it.
template <int NT, int OUT, int TO_SLOT>
__device__ __forceinline__ void drain_fast(uint32_t ta, const int32_t *s_gs, const float *s_sc, float w_scale, float *out0,
                                           float *const *outs, int n_out, int64_t idx0, int64_t ldo, int tvalid, bool nvalid,
                                           int32_t *slot_m, bool vec_ok = false, Scatter *sc = nullptr, int c_begin = 0,
                                           int c_end = NT) {
    // [c_begin, c_end): this warp's share of the NT columns (multiples of 32; two epilogue warpgroups split a tile)
    const bool ws = (w_scale != 1.0f);
    uint32_t va[16];
    ptx::tmem_ld_32x32b_x16(ta + c_begin, va);
    if constexpr (NT == 16) {
        tmem_ld_wait_dep(va);
        fast_chunk<OUT, TO_SLOT>(va, 0, s_gs, s_sc, ws, w_scale, out0, outs, n_out, idx0, ldo, tvalid, nvalid, slot_m, vec_ok, sc);
    } else {
        uint32_t vb[16];
#pragma unroll 1
        for (int c0 = c_begin; c0 < c_end; c0 += 32) {
            tmem_ld_wait_dep(va);
            ptx::tmem_ld_32x32b_x16(ta + c0 + 16, vb);
            fast_chunk<OUT, TO_SLOT>(va, c0, s_gs, s_sc, ws, w_scale, out0, outs, n_out, idx0, ldo, tvalid, nvalid, slot_m, vec_ok, sc);
            tmem_ld_wait_dep(vb);
            if (c0 + 32 < c_end) ptx::tmem_ld_32x32b_x16(ta + c0 + 32, va);
            fast_chunk<OUT, TO_SLOT>(vb, c0 + 16, s_gs, s_sc, ws, w_scale, out0, outs, n_out, idx0, ldo, tvalid, nvalid,
                                     slot_m, vec_ok, sc);
        }
    }
}

}  // namespace epi
}  // namespace vlut
