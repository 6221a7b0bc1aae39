// push.cu -- the all-gather of a row-sharded MUL_MAT DECOUPLED from the GEMM's accumulators (vlut_mul_mat_scatter_push).
//
// The fused scatter epilogue (epilogue.cuh, OUT = 1..3) writes every tile straight to the peers, so the drain of a tile
// runs at the NVLink's pace (13.9 us per 128 x 128 tile at N = 2) while it holds a TMEM accumulator: the MMAs of the tile
// after next wait for it and the step is GEMM + transfer almost in series (30 + 43 us at N = 2).  Here the GEMM stores its
// slice into the LOCAL result buffer only (full speed) and stamps a per-tile flag; this small kernel, launched on a second
// stream next to the GEMM (128 threads x 32 registers per CTA: it fits beside the GEMM's 480 x 128 on every SM), waits for
// each tile's flag and copies the tile from local memory (an L2 hit) to every peer while the GEMM goes on.
#include "vlut_common.cuh"

namespace vlut {

struct PushArgs {
    const float *src;     // this rank's [T][ldo] result buffer, at column col0
    float *dst[7];        // the peers' buffers, at column col0
    int n_dst;
    int64_t ldo, N, T;    // N = columns (rows of this rank's weight shard)
    int NT, TT, ntiles;   // token tile of the GEMM, token tiles, row-pair tiles * token tiles
    const uint32_t *tile_done;   // [ntiles][2]: stamped with `epoch` by the GEMM CTA (rank 0 / 1 of the pair) that finished it
    uint32_t epoch;
};

// (128 threads at <= 32 registers: 4096 registers, exactly what the CTA-pair kernel's 480 x 128 leave of an SM's file)
__global__ void __launch_bounds__(128, 16) push_kernel(const __grid_constant__ PushArgs a) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int items = 2 * a.ntiles;   // item = (tile, CTA of the pair): 128 rows x NT tokens
    const int i0 = (int) (((int64_t) blockIdx.x * items) / gridDim.x), i1 = (int) (((int64_t) (blockIdx.x + 1) * items) / gridDim.x);
    for (int it = i0; it < i1; it++) {
        const int tile = it >> 1, r = it & 1;
        if (threadIdx.x == 0) {
            uint32_t v;
            do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a.tile_done + it) : "memory");
                if (v != a.epoch) __nanosleep(200);
            } while (v != a.epoch);
        }
        __syncthreads();
        const int rp = tile / a.TT, tt = tile - rp * a.TT;
        const int64_t n0 = ((int64_t) 2 * rp + r) * TILE_ROWS;
        const int rows = (int) min((int64_t) TILE_ROWS, a.N - n0);
        const int64_t t0 = (int64_t) tt * a.NT;
        const int nt = (int) min((int64_t) a.NT, a.T - t0);
        if (rows <= 0 || 4 * lane >= rows) continue;   // (rows % 4 == 0: the caller checks N % 4 == 0)
        const float *s0 = a.src + t0 * a.ldo + n0 + 4 * lane;
        const int64_t off0 = t0 * a.ldo + n0 + 4 * lane;
        // one token row (<= 512 bytes) per warp instruction; four rows in flight per warp
        for (int t = warp * 4; t < nt; t += 16) {
            float4 v[4];
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (t + j < nt) v[j] = __ldcg(reinterpret_cast<const float4 *>(s0 + (int64_t) (t + j) * a.ldo));
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (t + j < nt) {
                    for (int d = 0; d < a.n_dst; d++)
                        *reinterpret_cast<float4 *>(a.dst[d] + off0 + (int64_t) (t + j) * a.ldo) = v[j];
                }
            }
        }
    }
}

cudaError_t launch_push(const float *src, float *const *dst, int n_dst, int64_t ldo, int64_t N, int64_t T, int NT, int TT,
                        int ntiles, const uint32_t *tile_done, uint32_t epoch, int sm_count, cudaStream_t s) {
    if (n_dst < 1 || n_dst > 7) return cudaErrorInvalidValue;
    PushArgs a{};
    a.src = src;
    for (int i = 0; i < n_dst; i++) a.dst[i] = dst[i];
    a.n_dst = n_dst;
    a.ldo = ldo; a.N = N; a.T = T;
    a.NT = NT; a.TT = TT; a.ntiles = ntiles;
    a.tile_done = tile_done;
    a.epoch = epoch;
    int grid = sm_count;
    if (grid > 2 * ntiles) grid = 2 * ntiles;
    push_kernel<<<grid, 128, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace vlut
