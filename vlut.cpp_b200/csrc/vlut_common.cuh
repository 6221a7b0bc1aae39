// vlut_common.cuh -- shared definitions of the sm_100a Vec-LUT MUL_MAT kernels.
//
// Device weight layout "T2" (what vlut_repack_weights produces and every GEMM kernel streams):
//   * 2-bit codes c = w + 1 in {0,1,2}  (same 2 bits/weight as GGUF I2_*; I1_* grows from 1.6 to 2 bits)
//   * matrix padded to Npad = ceil128(N) rows, Kpad = ceil128(K) columns, pad code = 1 (weight 0)
//   * tile  (rt, kb) = 128 rows x 128 k = 4096 B, tiles ordered [rt][kb]  -> one 128-row strip is contiguous
//   * tile  = 8 chunks of 16 rows, 512 B each
//   * chunk = 32 "lanes" x 16 B; lane = 4*g + tig (g = 0..7, tig = 0..3), 4 words of 32 bits per lane
//   * word  (rg, h) = 2*rg + h holds 16 consecutive k of ONE row:
//         row_in_chunk = 8*rg + g,   k_in_block = 32*tig + 16*h + 4*i + j
//     stored at byte j (0..3), bit pair i (0..3) of the word:  bits [8j+2i, 8j+2i+2)
//     so  (word >> 2i) & 0x03030303  is 4 consecutive k of that row as 4 u8 -- one mma.sync B register, or
//     one 32-bit quarter of a 16-byte run of the K-major tcgen05 A operand.
// A warp reading a chunk with one 128-bit load per lane gets, for mma.sync.m16n8k32 with the weight rows as
// the N dimension (B fragment: n = g, k-slot = 16*h + 4*tig + j), both B registers of k-step i for two
// 8-row groups.  The k-slot <-> k permutation inside a 128-block is matched by the activation fragments.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vlut_b200.h"

namespace vlut {

constexpr int TILE_ROWS = 128;
constexpr int TILE_K = 128;
constexpr int TILE_BYTES = TILE_ROWS * TILE_K / 4;  // 4096
constexpr int CHUNK_ROWS = 16;
constexpr int CHUNK_BYTES = 512;

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

// byte offset of the word holding (row n, k) in a T2 buffer, plus byte j / bit-pair i inside it
__host__ __device__ inline int64_t t2_word_offset(int64_t n, int64_t k, int64_t Kpad) {
    const int64_t rt = n / TILE_ROWS, kb = k / TILE_K;
    const int r = (int) (n % TILE_ROWS), kk = (int) (k % TILE_K);
    const int chunk = r / CHUNK_ROWS, rc = r % CHUNK_ROWS;
    const int rg = rc / 8, g = rc % 8;
    const int tig = kk / 32, h = (kk % 32) / 16;
    return (rt * (Kpad / TILE_K) + kb) * TILE_BYTES + chunk * CHUNK_BYTES + (4 * g + tig) * 16 + (2 * rg + h) * 4;
}
__host__ __device__ inline int t2_shift(int64_t k) {  // bit position of the code of k inside its word
    const int kk = (int) (k % 16);
    return 8 * (kk % 4) + 2 * (kk / 4);
}

// ---- activation workspace -------------------------------------------------------------------------------
// q_rm   int8  [Tpad][Kpad]      row-major (tcgen05 B operand via TMA, SIMPLE kernel), zero padded
// q_frag int8  [Kpad/128][Tpad/16][4 k-steps][32 lanes][16 B]   mma.sync A fragments, per lane
//              {a0,a1,a2,a3} = q[16mt+g][kq], q[16mt+g+8][kq], q[16mt+g][kq+16], q[16mt+g+8][kq+16],
//              kq = 128*kb + 32*tig + 4*i + (0..3)
// scales f32   [Tpad]
// gsum   int32 [Tpad][G]         sum_{k in group} q[t][k]   (the u8-code offset correction)
struct ActView {
    int8_t *q_rm;
    int8_t *q_frag;
    float *scales;
    int32_t *gsum;
    int64_t Kpad, Tpad;
    int G, gsize;
};

struct GemmArgs {
    const uint8_t *w;  // T2
    ActView act;
    float *out;
    int64_t ldo;
    int64_t K, N, T, Npad;
    int type;
    float w_scale;
    int32_t *part;      // int32 [Tpad_ws][Npad] stream-K partial sums (self-cleaning, zero between calls)
    int32_t *counters;  // per row-tile arrival counters (self-cleaning)
    int sm_count;
};

inline __device__ int32_t wrap16(int32_t v) { return (int32_t) (int16_t) (uint16_t) ((uint32_t) v & 0xffffu); }

// launchers (one per .cu)
cudaError_t launch_repack(int type, int64_t K, int64_t N, const uint8_t *gguf, uint8_t *t2, cudaStream_t s);
cudaError_t launch_unpack_t2(int64_t K, int64_t N, const uint8_t *t2, int8_t *w, cudaStream_t s);
cudaError_t launch_quantize(const float *x, int64_t ldx, int64_t K, int64_t T, const ActView &a, cudaStream_t s);
cudaError_t launch_gemm_simple(const GemmArgs &g, cudaStream_t s);
cudaError_t launch_gemm_mma(const GemmArgs &g, cudaStream_t s, int *launches);
bool gemm_mma_supported(int type, int64_t K, int64_t N, int64_t T);
cudaError_t launch_gemm_tc(const GemmArgs &g, cudaStream_t s, int *launches);
bool gemm_tc_supported(int type, int64_t K, int64_t N, int64_t T);
cudaError_t launch_gemm_lut(const GemmArgs &g, cudaStream_t s, int *launches);
bool gemm_lut_supported(int type, int64_t K, int64_t N, int64_t T);

}  // namespace vlut
