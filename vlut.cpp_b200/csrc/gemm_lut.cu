// gemm_lut.cu -- VLUT_ALGO_LUT: the faithful shared-memory vector-LUT formulation of the reference, on the GPU.
//
// This is the measured *candidate* north_star asks for (shared-memory 1->T vector-table lookup-accumulate), not the
// production path.  It restates the reference's algorithm (ggml/src/ggml-cpu/ggml-cpu-quants-vlut.c):
//   * gemm_make_table_* (:1873-2210): for every pack of 4 k build a table of all 81 ternary combinations,
//     table[idx][t] = sum_j w_j(idx) * q[t][k_j], one int16 lane per token;
//   * ADD_TABLE_ENTRIES_BLOCK* (:790-826): for every weight row, one packed byte -> one table row -> one vector add
//     into the row's running sums; int16 wrap per K-group (:80-87); fp32 epilogue (:1252-1261).
// On the GPU: a CTA owns 512 weight rows x one K-group (512 / 640 k).  Per 64-k chunk it builds 16 tables (one per byte
// position of the T2 row entry; index = the byte itself, 2-bit digits, 81 of 256 slots valid) of TP int16 lanes in
// shared memory, then every thread walks the 16 bytes of its two rows: one table row (TP x 2 bytes) per byte.
// Group partial sums are wrapped to int16 and combined with int32 atomics; a second tiny kernel applies the scale.
//
// Why it loses: per packed weight byte it reads 2*TP bytes of shared memory -- at the HBM rate
// of 23 B/clk/SM that is 46*TP B/clk against 128 B/clk of shared-memory bandwidth (with random-index bank conflicts on
// top), and the table build (81 x TP adds per 4 k) is amortised over only 512 rows instead of the CPU's N/nth.
#include "vlut_common.cuh"

namespace vlut {

constexpr int LUT_ROWS = 512;     // weight rows per CTA (4 strips)
constexpr int LUT_THREADS = 256;  // two rows per thread

template <int TP>
__global__ void __launch_bounds__(LUT_THREADS, 1) gemm_lut_kernel(GemmArgs p, int32_t *__restrict__ acc) {
    extern __shared__ __align__(16) uint8_t lut_smem[];
    int16_t *tab = reinterpret_cast<int16_t *>(lut_smem);                    // [16 tables][256 idx][TP]
    int8_t *q_s = reinterpret_cast<int8_t *>(lut_smem + 16 * 256 * TP * 2);  // [64 k][TP]
    const int tid = threadIdx.x;
    const int G = p.act.G;
    const int rb = blockIdx.x / G, g = blockIdx.x % G;  // row block (512 rows), K-group
    const int64_t Kpad = p.act.Kpad;
    const int64_t KC = Kpad / CHUNK_K;
    const int c_begin = g * (p.act.gsize / CHUNK_K);
    const int c_end = (int) min((int64_t) (c_begin + p.act.gsize / CHUNK_K), KC);

    int32_t sum[2][TP];
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
        for (int t = 0; t < TP; t++) sum[r][t] = 0;

    for (int kc = c_begin; kc < c_end; kc++) {
        __syncthreads();  // previous chunk's tables are no longer read
        // ---- stage q[t][64 k] transposed: q_s[k][t] ----
        for (int i = tid; i < CHUNK_K * TP; i += LUT_THREADS) {
            const int k = i / TP, t = i % TP;
            q_s[i] = (t < p.act.Tpad) ? p.act.q_rm[(int64_t) t * Kpad + (int64_t) kc * CHUNK_K + k] : (int8_t) 0;
        }
        __syncthreads();
        // ---- build the 16 tables: table b, entry idx (2-bit digits d_i = code of k = 16*j + 4*i + bb, b = 4*j + bb) ----
        for (int task = tid; task < 16 * 81; task += LUT_THREADS) {
            const int b = task / 81;
            int e = task % 81;
            int w[4], idx = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int d = e % 3;
                e /= 3;
                w[i] = d - 1;
                idx |= d << (2 * i);
            }
            const int j = b >> 2, bb = b & 3;
            int16_t *dst = tab + ((b * 256 + idx) * TP);
#pragma unroll
            for (int t = 0; t < TP; t++) {
                int v = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) v += w[i] * (int) q_s[(16 * j + 4 * i + bb) * TP + t];
                dst[t] = (int16_t) v;
            }
        }
        __syncthreads();
        // ---- lookup-accumulate: 16 packed bytes per row, one table row per byte ----
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int row = tid + r * LUT_THREADS;  // 0..511 inside the row block
            const int64_t rt = (int64_t) rb * (LUT_ROWS / TILE_ROWS) + row / TILE_ROWS;
            if (rt * TILE_ROWS >= p.Npad) continue;
            const uint4 wv = *reinterpret_cast<const uint4 *>(p.w + (rt * KC + kc) * TILE_BYTES + (row % TILE_ROWS) * 16);
            const uint32_t ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
            for (int b = 0; b < 16; b++) {
                const uint32_t idx = (ww[b >> 2] >> (8 * (b & 3))) & 0xffu;
                const int16_t *src = tab + ((b * 256 + idx) * TP);
#pragma unroll
                for (int t8 = 0; t8 < TP; t8 += 8) {
                    const uint4 v = *reinterpret_cast<const uint4 *>(src + t8);  // 8 int16 lanes
                    const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int h = 0; h < 4; h++) {
                        sum[r][t8 + 2 * h] += (int) (int16_t) (vv[h] & 0xffffu);
                        sum[r][t8 + 2 * h + 1] += (int) (int16_t) (vv[h] >> 16);
                    }
                }
            }
        }
    }
    // ---- K-group boundary: int16 wrap, combine across groups ----
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const int64_t n = (int64_t) rb * LUT_ROWS + tid + r * LUT_THREADS;
        if (n >= p.N) continue;
#pragma unroll
        for (int t = 0; t < TP; t++)
            if (t < p.T) {
                const int32_t v = wrap16(sum[r][t]);
                if (v != 0) atomicAdd(acc + (int64_t) t * p.Npad + n, v);
            }
    }
}

__global__ void gemm_lut_finish(GemmArgs p, int32_t *__restrict__ acc) {
    const int64_t n = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t t = blockIdx.y;
    if (n >= p.N) return;
    int32_t *a = acc + t * p.Npad + n;
    float v = (float) *a * p.act.scales[t];
    if (p.w_scale != 1.0f) v *= p.w_scale;
    p.out[t * p.ldo + n] = v;
    *a = 0;  // leave the accumulator clean for the next call
}

bool gemm_lut_supported(int type, int64_t K, int64_t N, int64_t T) {
    return type_valid(type) && T >= 1 && T <= 16 && K >= 4 && N >= 1;
}

// `g.part` is used as the int32 accumulator [16][Npad]; the caller guarantees it is zero on entry (and this kernel
// pair leaves it zero).
cudaError_t launch_gemm_lut(const GemmArgs &g, cudaStream_t s, int *launches) {
    if (g.n_out) return cudaErrorNotSupported;
    const int RB = (int) ceil_div(g.Npad, LUT_ROWS);
    const int grid = RB * g.act.G;
    int32_t *acc = g.part;
    cudaError_t e;
    if (g.T <= 8) {
        const size_t smem = 16 * 256 * 8 * 2 + 64 * 8;
        e = cudaFuncSetAttribute(gemm_lut_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (e != cudaSuccess) return e;
        gemm_lut_kernel<8><<<grid, LUT_THREADS, smem, s>>>(g, acc);
    } else {
        const size_t smem = 16 * 256 * 16 * 2 + 64 * 16;
        e = cudaFuncSetAttribute(gemm_lut_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (e != cudaSuccess) return e;
        gemm_lut_kernel<16><<<grid, LUT_THREADS, smem, s>>>(g, acc);
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    dim3 fg((unsigned) ceil_div(g.N, 256), (unsigned) g.T);
    gemm_lut_finish<<<fg, 256, 0, s>>>(g, acc);
    if (launches) *launches += 2;
    return cudaGetLastError();
}

}  // namespace vlut
