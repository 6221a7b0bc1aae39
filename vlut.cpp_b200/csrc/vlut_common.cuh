// vlut_common.cuh -- shared definitions of the sm_100a Vec-LUT MUL_MAT kernels.
//
// Device weight layout "T2" (what vlut_repack_weights produces and every GEMM kernel streams):
//   * 2-bit codes c = w + 1 in {0,1,2}  (same 2 bits/weight as GGUF I2_*; I1_* grows from 1.6 to 2 bits)
//   * matrix padded to Npad = ceil128(N) rows, Kpad = ceil128(K) columns, pad code = 1 (weight 0)
//   * strip rt = 128 consecutive rows; inside a strip the K axis is cut into 64-wide chunks kc; one
//     (strip, chunk) tile is 128 rows x 16 bytes = 2048 B, row-major:   tile[row][16 B]
//     tiles ordered [rt][kc]  ->  a strip (and any K-group of it) is one contiguous run of bytes (TMA bulk)
//   * the 16 bytes of a row are 4 little-endian 32-bit words; word j holds 16 consecutive k of that row,
//         k_in_chunk = 16*j + 4*i + b     stored at byte b (0..3), bit pair i (0..3):  bits [8b+2i, 8b+2i+2)
//     so  (word >> 2i) & 0x03030303  is 4 consecutive k of the row as 4 u8 codes: one 32-bit column of the
//     K-major tcgen05 A operand in TMEM (lane = row, column = k/4), or a quarter of a 16-byte swizzle chunk of
//     the shared-memory A operand.  Thread m of a 128-thread decode group owns row m: its 128-bit shared-memory
//     load is conflict-free and yields two K=32 MMA steps.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vlut_b200.h"

namespace vlut {

constexpr int TILE_ROWS = 128;
constexpr int CHUNK_K = 64;                       // k per tile
constexpr int TILE_BYTES = TILE_ROWS * 16;        // 2048
constexpr int BLOCK_K = 128;                      // k per pipeline stage (2 tiles)
constexpr int BLOCK_BYTES = 2 * TILE_BYTES;       // 4096

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ inline int64_t round_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }

__host__ __device__ inline bool type_is_i1(int type) { return type == VLUT_TYPE_I1_V || type == VLUT_TYPE_I1_V_2; }
__host__ __device__ inline bool type_is_i2(int type) {
    return type == VLUT_TYPE_I2_V || type == VLUT_TYPE_I2_V_4 || type == VLUT_TYPE_I2_V_8;
}
__host__ __device__ inline bool type_valid(int type) { return type_is_i1(type) || type_is_i2(type); }
// K-group inside which the reference keeps int16 partial sums (ggml-cpu-quants-vlut.c:1191, :1540)
__host__ __device__ inline int group_size(int type) { return type_is_i1(type) ? 640 : 512; }
// byte interleave of the GGUF packing (type_traits_vlut.tile_size, ggml-cpu.c:1180-1238)
__host__ __device__ inline int gguf_interleave(int type) {
    return type == VLUT_TYPE_I2_V_4 ? 4 : type == VLUT_TYPE_I2_V_8 ? 8 : type == VLUT_TYPE_I1_V_2 ? 2 : 1;
}
__host__ __device__ inline int64_t gguf_row_bytes(int type, int64_t K) {
    return type_is_i1(type) ? (K / 20 * 4 + K % 20 / 4) : K / 4;
}

// byte offset of the word holding (row n, k) in a T2 buffer, and the bit position of k's code inside it
__host__ __device__ inline int64_t t2_word_offset(int64_t n, int64_t k, int64_t Kpad) {
    const int64_t rt = n / TILE_ROWS, kc = k / CHUNK_K;
    const int r = (int) (n % TILE_ROWS), j = (int) (k % CHUNK_K) / 16;
    return (rt * (Kpad / CHUNK_K) + kc) * TILE_BYTES + r * 16 + j * 4;
}
__host__ __device__ inline int t2_shift(int64_t k) {
    const int kk = (int) (k % 16);
    return 8 * (kk % 4) + 2 * (kk / 4);
}

// token tile (UMMA N) used for a token count
__host__ __device__ inline int token_tile(int64_t T) { return T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : 128; }
// The CTA-pair kernel also has a 256-token tile (one decoded A stage feeds two 128-token halves: weights unpacked T/256
// instead of T/128 times, 2048 clk of MMA per hand-off, but a single accumulator, so every tile pays an un-overlapped drain).
// Which tile is faster depends on K (the drain bubble amortises over K), on how the tiles quantise into rounds on the 74 CTA
// pairs, and on the type (I1: 640-k groups).  Cost model in SM clocks, fitted to measurements at four shapes x six token
// counts (profiles/r2_tcp_tiles.txt: it picks the faster tile in 23 of 24 cases, the miss costs 6 %):
//     tile time = G * unit + tile_overhead;  whole / hybrid schedule as in launch_out (gemm_tc5.cu);  + fixed
inline double pair_tile_cost(int64_t K, int64_t N, int64_t T, int NT, int gsize, int sm_count) {
    // constants in clk per 512 k of a tile, fitted to scripts/time_tcp.py on the lean-loop kernel (profiles/r2_tcp_tiles_lean.txt:
    // the model picks the faster tile in 24 of 24 measured cases).  Shapes whose K-groups do not fit the lean loops (640-k
    // groups, K not a multiple of the group size) walk K in uniform blocks on the fast path ("flat", gemm_tc5.cuh): the
    // same constants apply to ceil(K / 512) blocks.
    const double unit = NT > 128 ? 1900.0 : 1300.0;
    const double tile = NT > 128 ? 6000.0 : 0.0, fixed = NT > 128 ? 9000.0 : 6000.0;
    (void) gsize;
    const int64_t G = ceil_div(K, 512), RP = (ceil_div(N, TILE_ROWS) + 1) / 2, TT = ceil_div(T, NT);
    const int64_t ptiles = RP * TT, npairs = sm_count / 2 > 0 ? sm_count / 2 : 1;
    const int64_t rounds = ceil_div(ptiles, npairs), full = ptiles / npairs, su = (ptiles - full * npairs) * G;
    const double whole = (double) rounds * (G * unit + tile);
    const double hybrid = (double) full * (G * unit + tile) + (su ? (1.3 * ceil_div(su, npairs) + 3.0) * unit : 0.0);
    return (full >= 1 && hybrid < whole ? hybrid : whole) + fixed;
}
inline int token_tile_pair(int64_t K, int64_t N, int64_t T, int gsize, int sm_count, int forced) {
    if (forced == 64 || forced == 128 || forced == 256) return (forced == 64 && T > 64) ? 128 : forced;
    if (T <= 64) return 64;
    if (T <= 128) return 128;
    return pair_tile_cost(K, N, T, 256, gsize, sm_count) < pair_tile_cost(K, N, T, 128, gsize, sm_count) ? 256 : 128;
}

// ---- activation workspace -------------------------------------------------------------------------------
// q_rm   int8  [Tpad][Kpad]   row-major, zero padded (SIMPLE kernel, test read-back)
// q_sw   int8  [Tpad/NT][Kpad/128][NT][128 B]   K-major UMMA operand tiles, already 128B-swizzled: the 16-byte
//              chunk c of token row t sits at chunk position c ^ (t % 8).  One [NT][128 B] tile dropped into shared
//              memory by a bulk copy IS a SWIZZLE_128B K-major [NT x 128] s8 operand (8-row atoms of 1024 B); the
//              tiles of consecutive 128-k sub-blocks of one token tile are contiguous (one copy per stage).
// scales f32   [Tpad]
// gpre   int32 [Tpad][G+1]    exclusive prefix sums over the K-groups of sum_{k in group} q[t][k]: the code-offset
//                             correction of groups [a, b] is gpre[b+1] - gpre[a]  (undoes the +1 of the u8 codes)
// nowrap int32 [Tpad]         1 if for every group of token t sum_k |q| <= 32767, i.e. no int16 wrap is possible
//                             whatever the weights are (full-K accumulation is then exact)
// tile_unsafe u32 [Tpad/16+1] [tt] = `epoch` of this quantisation iff some token of tile tt can wrap; [Tpad/16] = `epoch` iff any tile can (any unsafe token's CTA
//                             stores the epoch: no reduction, no atomics); epoch = per-context call counter, never 0
struct ActView {
    int8_t *q_rm;
    int8_t *q_sw;
    float *scales;
    int32_t *gpre;
    int32_t *nowrap;
    uint32_t *tile_unsafe;
    uint32_t epoch;
    int64_t Kpad, Tpad;
    int G, gsize, NT;
    int sw_tile;  // tokens per swizzled tile of q_sw: NT, or NT / 2 for the CTA-pair kernel (each CTA stages one half tile)
    // K order inside every 128-k sub-block row of q_sw.  0: natural.  1 (small-T kernel, mask-only weight unpack): the
    // 128 bytes are regrouped by bit-pair position i of the packed weight words: byte i*32 + w*4 + b holds k = 16 w + 4 i + b
    // (w = 0..7 the 16-k word group, b = 0..3), so that the 32 bytes of one K=32 MMA all pair with weight codes that sit
    // at the same bit position (i) of their words -- the unpack is then a mask, the shift is folded into 4 accumulators.
    int perm;
};

// byte position of k (0..127, inside a 128-k sub-block) in a q_sw row
__host__ __device__ inline int act_row_pos(int k128, int perm) {
    return perm ? ((k128 & 15) >> 2) * 32 + (k128 >> 4) * 4 + (k128 & 3) : k128;
}

// one weight of a batched launch (several weights that share the quantised activations)
constexpr int MAX_BATCH = 4;
struct TensorRef {
    const uint8_t *w;   // T2
    float *out;         // f32 [T][ldo]
    int64_t N, ldo;
    float w_scale;
    int strip0;         // first 128-row strip of this tensor in the launch's concatenated strip index space
};

struct GemmArgs {
    TensorRef tens[MAX_BATCH + 1];   // tens[n_tens].strip0 = total strips (sentinel); tens[0] mirrors the fields below
    int n_tens;
    const uint8_t *w;  // T2
    ActView act;
    float *out;         // single-GPU destination, f32 [T][ldo]
    // multi-GPU fused all-gather (row-sharded weights): the epilogue writes its column slice of the [T][N_total]
    // result into EVERY rank's buffer -- either n_out peer pointers (CUDA P2P over NVLink), or one NVSwitch multicast
    // address (multimem.st: the switch replicates the store to all ranks).  Pointers are already offset to this
    // rank's first column.  n_out == 0: plain single-destination store through `out`.
    float *outs[8];
    int n_out;
    int multicast;
    int vec_ok;         // scatter destinations are 16-byte aligned with ldo % 4 == 0: four rows per lane per store
    int bulk_ok;        // ... and N % 4 == 0, peer pointers: the CTA-pair kernel stages tiles in shared memory and scatters
                        // them with bulk async copies (epilogue.cuh, Scatter)
    uint32_t *tile_done;   // push mode (push.cu): [tiles][2] stamped with tile_epoch when a CTA's 128 x NT block of a tile is in `out`
    uint32_t tile_epoch;
    int local_bulk;     // n_out == 0 and `out` allows the staged drain (16-byte aligned, ldo % 4 == 0, N % 4 == 0): OUT = 4
    int64_t ldo;
    int64_t K, N, T, Npad;
    int type;
    float w_scale;
    int32_t *part;      // int32 [CTAs][4 slots + 2 scratch][NT][128] stream-K partial tiles (no initialisation needed)
    int32_t *counters;  // per output-tile arrival counters (self-cleaning)
    int32_t *acc;       // int32 [CTAs][128][128] accumulation tiles of split output tiles (TCP): contributors red.add their
                        // partials, the last arriver reads the sum and writes zeros back (zeroed at allocation, self-cleaning)
    int sm_count;
    int device;         // CUDA device ordinal (per-device caches of function attributes in the launchers)
};

// epilogue store, selected at compile time: 0 local, 1 peer scatter (NVLink P2P), 2 NVSwitch multicast (multimem.st),
// 3 peer scatter through shared-memory staging + bulk async copies (these direct stores are then only the exact path's),
// 4 the same staging for the single local destination
template <int MODE>
__device__ __forceinline__ void store_out(float *const out, float *const *outs, int n_out, int64_t idx, float v) {
    if (MODE == 0) {
        out[idx] = v;
    } else if (MODE == 2) {
        asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(out + idx), "f"(v) : "memory");
    } else {
        for (int i = 0; i < n_out; i++) outs[i][idx] = v;
    }
}

// 16-byte flavours of the scatter stores: four consecutive rows n .. n+3 of one token
template <int MODE>
__device__ __forceinline__ void store_out4(float *const out, float *const *outs, int n_out, int64_t idx, float4 v) {
    if (MODE == 0) {
        *reinterpret_cast<float4 *>(out + idx) = v;
    } else if (MODE == 2) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(out + idx), "f"(v.x), "f"(v.y),
                     "f"(v.z), "f"(v.w)
                     : "memory");
    } else {
        for (int i = 0; i < n_out; i++) *reinterpret_cast<float4 *>(outs[i] + idx) = v;
    }
}

inline __device__ int32_t wrap16(int32_t v) { return (int32_t) (int16_t) (uint16_t) ((uint32_t) v & 0xffffu); }

// ---- run-time knobs ------------------------------------------------------------------------------------------------
// Every VLUT_* environment variable the library understands, read ONCE (first use, vlut_ctx_create, or the test hook
// vlut_reload_env) -- never on the launch path.  -1 = not set (the launcher's own policy applies).
struct Knobs {
    int algo;                      // VLUT_B200_ALGO      force a GEMM algorithm (VLUT_ALGO_*)
    int host_chunk;                // VLUT_HOST_CHUNK     token chunk of the host-buffer pipeline (0 = no pipeline)
    int pdl;                       // VLUT_PDL            programmatic dependent launch (default 1)
    int tc_fast;                   // VLUT_TC_FAST        0 forces the exact (per-K-group drain) path
    int tc_smem_kb, tc_stages;     // VLUT_TC_SMEM_KB, VLUT_TC_STAGES
    int tc_coop_timeout;                    // VLUT_TC_COOP_TIMEOUT: clk a contributor waits for the others before the tile falls back to the last-arriver fix-up (default 400000)
    int tc_coop;                            // VLUT_TC_COOP: contributors of a hybrid split tile share its fix-up (default 1)
    int tc_grid, tc_whole, tc_split_last;   // VLUT_TC_GRID, VLUT_TC_WHOLE, VLUT_TC_SPLIT_LAST (tests: force the split paths)
    int tc_abl, tc_debug, tc_debug_skip;    // VLUT_TC_ABL, VLUT_TC_DEBUG, VLUT_TC_DEBUG_SKIP
    int tc_nt;                     // VLUT_TC_NT          force the TCP token tile (128 / 256)
    int tc_flat;                   // VLUT_TC_FLAT        0: never walk K in uniform A blocks across the K-group boundaries (comparison, tests)
    int tc_lean;                   // VLUT_TC_LEAN        0: the generic (unit-walking) decode loop even for uniform shapes (comparison, tests)
    int tc_astages;                // VLUT_TC_ASTAGES     2: two A stages of 512 k, 4: four of 256 k; default per token tile
    int tcs_nas, tcs_split, tcs_issuers, tcs_depth, tcs_dbg_cta, tcs_dbg_mask, tcs_pre;   // VLUT_TCS_*
    int tcs_rows;                  // VLUT_TCS_ROWS       force the rows per (virtual) strip of the small-T kernel (64..128, multiple of 4; 128 = off)
    int quant_kernel, quant_threads, quant_pdl, quant_early, quant_tpc;   // VLUT_QUANT_*
    int gemv_split, gemv_max_t;    // VLUT_GEMV_SPLIT, VLUT_GEMV_MAX_T
    int scatter_vec;               // VLUT_SCATTER_VEC    0: scalar 4-byte scatter stores (comparison)
    int tc_vec_out;                // VLUT_TC_VEC_OUT     1: CTA-pair kernel, single-GPU output: 16-byte stores after a 4 x 4 lane transpose (measured 8-25 % SLOWER than the 4-byte stores, profiles/r2_tcp_drain_ablation.txt; default 0)
    int tc_bulk_out;               // VLUT_TC_BULK_OUT    1: the CTA-pair kernel drains single-GPU outputs through shared memory + one 2-D TMA tensor store per 16-token chunk (measured: T = 512 27.4 vs 25.6 us direct, equal at T = 2048; sixteen 512-byte bulk copies per chunk: 31-42 us)
    int scatter_bulk;              // VLUT_SCATTER_BULK   0: no staged bulk-copy scatter in the CTA-pair kernel (comparison)
};
const Knobs &knobs();
void knobs_reload();

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel instantiation, device) instead of on every launch:
// `cache` is a static int[MAX_DEVICES] owned by the calling launcher template
constexpr int MAX_DEVICES = 16;
template <typename Kern>
inline cudaError_t ensure_dyn_smem(Kern kern, size_t smem, int device, int *cache) {
    if (device >= 0 && device < MAX_DEVICES && cache[device] >= (int) smem) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e == cudaSuccess && device >= 0 && device < MAX_DEVICES) cache[device] = (int) smem;
    return e;
}

long long *debug_buffer(cudaStream_t s, int skip);   // c_api.cu: zeroed buffer of the kernels' optional instrumentation

cudaError_t measure_int8_peak(int sm_count, cudaStream_t s, int iters, double *tops, double *clk_per_mma);  // peak.cu

cudaError_t launch_barrier(uint32_t *const *flags, int n, int rank, cudaStream_t s);   // barrier.cu
cudaError_t launch_push(const float *src, float *const *dst, int n_dst, int64_t ldo, int64_t N, int64_t T, int NT, int TT,
                        int ntiles, const uint32_t *tile_done, uint32_t epoch, int sm_count, cudaStream_t s);   // push.cu
// launchers (one per .cu)
cudaError_t launch_repack(int type, int64_t K, int64_t N, const uint8_t *gguf, uint8_t *t2, cudaStream_t s);
cudaError_t launch_unpack_t2(int64_t K, int64_t N, const uint8_t *t2, int8_t *w, cudaStream_t s);
// pdl: launch with the programmatic-serialisation attribute; early: row loads ahead of griddepcontrol.wait (the host has
// proved that no kernel in flight writes x); want_rm: also write the row-major copy q_rm
cudaError_t launch_quantize(const float *x, int64_t ldx, int64_t K, int64_t T, const ActView &a, cudaStream_t s, bool pdl,
                            bool early, bool want_rm);
cudaError_t launch_relayout(const ActView &a, cudaStream_t s);   // rewrite q_sw from q_rm for a.sw_tile / a.perm
cudaError_t launch_unswizzle(const ActView &a, cudaStream_t s);  // materialise q_rm from q_sw
cudaError_t launch_gemm_simple(const GemmArgs &g, cudaStream_t s);
cudaError_t launch_scale_rows(const float *x, int64_t ldx, float *out, int64_t ldo, int64_t N, int64_t T, float sc,
                              cudaStream_t s);
cudaError_t launch_gemm_tc5(const GemmArgs &g, cudaStream_t s, int *launches);  // CTA pairs, cta_group::2
bool gemm_tc5_supported(int type, int64_t K, int64_t N, int64_t T);
cudaError_t launch_gemm_tcs(const GemmArgs &g, cudaStream_t s, int *launches);  // small T: cp.async weight stream, cluster split-K
bool gemm_tcs_supported(int type, int64_t K, int64_t N, int64_t T);
// whole op in one launch (activation quantisation fused), T <= 8
cudaError_t launch_gemm_gemv(const GemmArgs &g, const float *x, int64_t ldx, cudaStream_t s, int *launches);
bool gemm_gemv_supported(int type, int64_t K, int64_t N, int64_t T, const float *x, int64_t ldx);
cudaError_t launch_gemm_lut(const GemmArgs &g, cudaStream_t s, int *launches);
bool gemm_lut_supported(int type, int64_t K, int64_t N, int64_t T);

}  // namespace vlut
