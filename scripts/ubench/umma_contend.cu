// umma_contend.cu -- micro-benchmark: what slows a chain of small tcgen05.mma (kind::i8, M=128, N=16, A in TMEM) down
// inside the small-T GEMM kernels?  One CTA per SM; warp 0 issues MMAs (K=32 each, 8 per "block", accumulating into one
// TMEM accumulator); optional company:
//   bit 0  four warps spin on an mbarrier (mbarrier.try_wait) like idle epilogue / producer warps do
//   bit 1  four warps stream tcgen05.st.x16 into other TMEM columns (the decode warps), with wait::st per 2 stores
//   bit 2  a tcgen05.commit after every 8 MMAs (the A-stage release)
//   bit 3  the issuing warp re-elects (elect.sync + __syncwarp) for every 4 MMAs, as the kernel's per-sub-block loop does
//   bit 4  B descriptor and A column change from MMA to MMA (walk over 8 KB of shared memory / 64 columns)
// Prints cycles per MMA (issue, and to completion).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../vlut.cpp_b200/csrc/ptx.cuh"
using namespace vlut;

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

template <int N>
__global__ void __launch_bounds__(288, 1) k(long long *out, int blocks, int mode) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar_done, bar_dummy, bar_never;
    __shared__ uint32_t tbase;
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) ((uint32_t *) smem_raw)[i] = 0x01010101u;
    if (threadIdx.x == 0) {
        ptx::mbar_init(ptx::smem_u32(&bar_done), 1);
        ptx::mbar_init(ptx::smem_u32(&bar_dummy), 1);
        ptx::mbar_init(ptx::smem_u32(&bar_never), 1);
        ptx::fence_barrier_init();
    }
    if (warp == 0) { ptx::tmem_alloc(ptx::smem_u32(&tbase), 512); ptx::tmem_relinquish(); }
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = tbase;
    if (warp == 0) {
        constexpr uint32_t idesc = make_idesc(N);
        long long t0 = clock64(), t1 = 0, t2 = 0;
        for (int b = 0; b < blocks; b++) {
            if (mode & 8) {
                for (int s = 0; s < 2; s++) {
                    if (ptx::elect_one()) {
#pragma unroll
                        for (int ks = 0; ks < 4; ks++) {
                            const int i = (mode & 16) ? (b * 8 + s * 4 + ks) : ks;
                            ptx::umma_i8_ts(tm, tm + 256 + (i % 8) * 8, make_smem_desc(smem_base + ((i / 4) % 4) * 2048) + (i % 4) * 2,
                                            idesc, 1);
                        }
                    }
                    __syncwarp();
                }
                if ((mode & 4) && ptx::elect_one()) ptx::umma_commit(ptx::smem_u32(&bar_dummy));
                __syncwarp();
            } else if (ptx::elect_one()) {
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int i = (mode & 16) ? (b * 8 + j) : (j % 4);
                    ptx::umma_i8_ts(tm, tm + 256 + (i % 8) * 8, make_smem_desc(smem_base + ((i / 4) % 4) * 2048) + (i % 4) * 2, idesc, 1);
                }
                if (mode & 4) ptx::umma_commit(ptx::smem_u32(&bar_dummy));
            }
            __syncwarp();
        }
        t1 = clock64();
        if (ptx::elect_one()) ptx::umma_commit(ptx::smem_u32(&bar_done));
        __syncwarp();
        ptx::mbar_wait(ptx::smem_u32(&bar_done), 0);
        t2 = clock64();
        if (lane == 0) { out[blockIdx.x * 2] = t1 - t0; out[blockIdx.x * 2 + 1] = t2 - t0; }
    } else if (warp >= 1 && warp <= 4) {
        if (mode & 1) {
            // idle pollers: wait for the end of the run on bar_done (one try_wait loop, as the kernels' mbar_wait)
            ptx::mbar_wait(ptx::smem_u32(&bar_done), 0);
        }
    } else if (warp >= 5 && warp <= 8) {
        if (mode & 2) {
            const int q = warp & 3;   // TMEM lane quadrant of this warp
            uint32_t r[16];
#pragma unroll
            for (int i = 0; i < 16; i++) r[i] = 0x01010101u * (i & 3);
            const uint32_t ta = tm + ((uint32_t) (q * 32) << 16) + 320;
            while (!ptx::mbar_try_wait(ptx::smem_u32(&bar_done), 0)) {
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    ptx::tmem_st_32x32b_x16(ta + c * 16, r);
                    ptx::tmem_st_32x32b_x16(ta + 64 + c * 16, r);
                    ptx::tmem_st_wait();
                }
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tm, 512); }
}

int main() {
    long long *d;
    cudaMalloc(&d, 148 * 2 * sizeof(long long));
    const int blocks = 128;
    auto kern = k<16>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const int modes[] = {0, 1, 2, 4, 8, 16, 3, 12, 28, 31, 29, 30};
    for (int mode : modes) {
        kern<<<148, 288, 64 * 1024>>>(d, blocks, mode);
        kern<<<148, 288, 64 * 1024>>>(d, blocks, mode);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[296];
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        double issue = 0, total = 0;
        for (int i = 0; i < 148; i++) { issue += h[2 * i]; total += h[2 * i + 1]; }
        issue /= 148; total /= 148;
        printf("mode %2d [%s%s%s%s%s]: issue %.1f clk/MMA, complete %.1f clk/MMA  (%s)\n", mode, (mode & 1) ? "spin " : "",
               (mode & 2) ? "st " : "", (mode & 4) ? "commit " : "", (mode & 8) ? "re-elect " : "", (mode & 16) ? "walk " : "",
               issue / (blocks * 8), total / (blocks * 8), cudaGetErrorString(e));
        fflush(stdout);
    }
    cudaFree(d);
    return 0;
}
