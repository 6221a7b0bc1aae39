// red_rate.cu -- what does it cost to park a 128 x 128 int32 partial tile in an L2 accumulation tile (split output tiles of the
// CTA-pair kernel, gemm_tc5.cuh)?  148 CTAs x 128 threads, each CTA adds TILES partial tiles of 128 columns x 128 rows into its
// own 64 KB slot, all CTAs at once:
//   mode 0  red.global.add.s32, one per value, a warp covers 128 contiguous bytes (the shipped layout: slot[c * 128 + m])
//   mode 1  red.global.add.v4.f32: four columns per instruction, a warp covers 512 contiguous bytes (exact for |sum| < 2^24)
//   mode 2  red.global.add.v2.f32
//   mode 3  st.global.v4 (no accumulation: the floor of moving the bytes)
//   mode 4  red.global.add.s32 with the v4 layout (lane stride 16 B: 4 instructions cover the same 512 B)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_rate red_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(128) k(int32_t *acc, int tiles, long long *clk) {
    int32_t *slot = acc + (size_t) blockIdx.x * 128 * 128;
    const int m = threadIdx.x;
    const long long t0 = clock64();
    for (int t = 0; t < tiles; t++) {
#pragma unroll 4
        for (int c = 0; c < 128; c += 4) {
            const int v0 = c + t + m, v1 = v0 + 1, v2 = v0 + 2, v3 = v0 + 3;
            if (MODE == 0) {
                asm volatile("red.global.add.s32 [%0], %1;" ::"l"(slot + (c + 0) * 128 + m), "r"(v0) : "memory");
                asm volatile("red.global.add.s32 [%0], %1;" ::"l"(slot + (c + 1) * 128 + m), "r"(v1) : "memory");
                asm volatile("red.global.add.s32 [%0], %1;" ::"l"(slot + (c + 2) * 128 + m), "r"(v2) : "memory");
                asm volatile("red.global.add.s32 [%0], %1;" ::"l"(slot + (c + 3) * 128 + m), "r"(v3) : "memory");
            } else if (MODE == 1) {
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(slot + (c / 4 * 128 + m) * 4), "f"((float) v0),
                             "f"((float) v1), "f"((float) v2), "f"((float) v3)
                             : "memory");
            } else if (MODE == 2) {
                asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(slot + (c / 4 * 128 + m) * 4), "f"((float) v0), "f"((float) v1) : "memory");
                asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(slot + (c / 4 * 128 + m) * 4 + 2), "f"((float) v2), "f"((float) v3) : "memory");
            } else if (MODE == 3) {
                asm volatile("st.global.v4.s32 [%0], {%1, %2, %3, %4};" ::"l"(slot + (c / 4 * 128 + m) * 4), "r"(v0), "r"(v1), "r"(v2), "r"(v3) : "memory");
            } else {
                asm volatile("red.global.add.s32 [%0], %1;" ::"l"(slot + (c / 4 * 128 + m) * 4 + 0), "r"(v0) : "memory");
                asm volatile("red.global.add.s32 [%0], %1;" ::"l"(slot + (c / 4 * 128 + m) * 4 + 1), "r"(v1) : "memory");
                asm volatile("red.global.add.s32 [%0], %1;" ::"l"(slot + (c / 4 * 128 + m) * 4 + 2), "r"(v2) : "memory");
                asm volatile("red.global.add.s32 [%0], %1;" ::"l"(slot + (c / 4 * 128 + m) * 4 + 3), "r"(v3) : "memory");
            }
        }
    }
    __threadfence();
    __syncthreads();
    if (m == 0) clk[blockIdx.x] = clock64() - t0;
}

template <int MODE>
void run(const char *name, int grid, int tiles, int32_t *acc, long long *clk) {
    cudaMemset(acc, 0, (size_t) 148 * 128 * 128 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<grid, 128>>>(acc, tiles, clk);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<grid, 128>>>(acc, tiles, clk);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[148]; cudaMemcpy(h, clk, grid * sizeof(long long), cudaMemcpyDeviceToHost);
    double mean = 0; long long mx = 0;
    for (int i = 0; i < grid; i++) { mean += h[i]; mx = h[i] > mx ? h[i] : mx; }
    mean /= grid;
    printf("%-34s grid %3d, %2d tiles per CTA: %7.2f us kernel, %6.0f clk per tile per CTA (mean; slowest CTA %6.0f)  %s\n", name, grid, tiles,
           ms * 1e3, mean / tiles, (double) mx / tiles, cudaGetErrorString(e));
}

int main() {
    int32_t *acc; long long *clk;
    cudaMalloc(&acc, (size_t) 148 * 128 * 128 * 4); cudaMalloc(&clk, 148 * sizeof(long long));
    for (int grid : {148, 8}) {
        for (int tiles : {1, 8}) {
            run<0>("red.add.s32 (shipped layout)", grid, tiles, acc, clk);
            run<4>("red.add.s32 (16 B lane stride)", grid, tiles, acc, clk);
            run<2>("red.add.v2.f32", grid, tiles, acc, clk);
            run<1>("red.add.v4.f32", grid, tiles, acc, clk);
            run<3>("st.v4 (no accumulation)", grid, tiles, acc, clk);
        }
    }
    return 0;
}
