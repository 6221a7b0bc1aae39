// pair_mma_rate.cu -- throughput of tcgen05.mma.cta_group::2 kind::i8 (M = 256 over a CTA pair, A in TMEM, B half in each
// CTA's shared memory) and what a tcgen05.commit (multicast to both CTAs) every C MMAs costs the issuing thread and the pipe.
//   one cluster per SM pair, one issuing thread in the even CTA; all 74 pairs run.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pair_mma_rate pair_mma_rate.cu
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
    return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (256 >> 4) << 24);
}

// C: MMAs between commits (0 = no commit inside the loop); NB: distinct B tiles cycled through (shared-memory footprint)
template <int N, int C, int DBL>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) k(long long *out, int iters) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar, sink[4];
    __shared__ uint32_t tbase;
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t rank = ptx::cluster_ctarank();
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += 128) ((uint32_t *) smem_raw)[i] = 0x01010101u;
    if (threadIdx.x == 0) {
        ptx::mbar_init(ptx::smem_u32(&bar), 1);
        for (int i = 0; i < 4; i++) ptx::mbar_init(ptx::smem_u32(&sink[i]), 1);
        ptx::fence_barrier_init();
    }
    if (threadIdx.x < 32) { ptx::tmem_alloc_pair(ptx::smem_u32(&tbase), 512); ptx::tmem_relinquish_pair(); }
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tm = tbase;
    if (threadIdx.x < 32 && rank == 0) {
        constexpr uint32_t idesc = make_idesc(N);
        long long t0 = 0, t1 = 0, t2 = 0;
        if (ptx::elect_one()) {
            t0 = clock64();
            int since = 0, sk = 0;
            for (int it = 0; it < iters; it++) {
                const uint64_t b_desc = make_smem_desc(smem_base + (it & 3) * 16384);
#pragma unroll
                for (int ks = 0; ks < 4; ks++) ptx::umma_i8_ts_pair(tm, tm + 256 + (it & 7) * 32 + ks * 8, b_desc + ks * 2, idesc, 1);
                if (C > 0) {
                    since += 4;
                    if (since >= C) {
                        since = 0;
                        ptx::umma_commit_pair(ptx::smem_u32(&sink[sk]), 3);
                        sk = (sk + 1) & 3;
                        if (DBL) {   // a second commit straight after the first (stage-empty + A-empty in the GEMM kernel)
                            ptx::umma_commit_pair(ptx::smem_u32(&sink[sk]), 3);
                            sk = (sk + 1) & 3;
                        }
                    }
                }
            }
            t1 = clock64();
            ptx::umma_commit_pair(ptx::smem_u32(&bar), 3);
        }
        __syncwarp();
        ptx::mbar_wait(ptx::smem_u32(&bar), 0);
        t2 = clock64();
        if (threadIdx.x == 0) { out[blockIdx.x] = t1 - t0; out[blockIdx.x + 1] = t2 - t0; }
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc_pair(tm, 512); }
}

template <int N, int C, int DBL = 0>
void run() {
    long long *d; cudaMalloc(&d, 148 * sizeof(long long));
    const int iters = 512;
    auto kern = k<N, C, DBL>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    for (int grid : {2, 148}) {
        kern<<<grid, 128, 100 * 1024>>>(d, iters);
        kern<<<grid, 128, 100 * 1024>>>(d, iters);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[148]; cudaMemcpy(h, d, grid * sizeof(long long), cudaMemcpyDeviceToHost);
        double issue = 0, total = 0; for (int i = 0; i < grid; i += 2) { issue += h[i]; total += h[i + 1]; }
        issue /= grid / 2; total /= grid / 2;
        printf("pair i8 TS N=%3d %s commit every %2d MMAs, grid %3d: issue %.1f clk/MMA, complete %.1f clk/MMA (nominal %d)  (%s)\n", N, DBL ? "DOUBLE" : "single", C, grid,
               issue / (iters * 4), total / (iters * 4), N / 2, cudaGetErrorString(e));
    }
    cudaFree(d);
}

int main() {
    run<128, 0>(); run<128, 16>(); run<128, 8>(); run<128, 4>();
    run<128, 16, 1>(); run<128, 8, 1>(); run<128, 4, 1>();
    run<256, 0>(); run<256, 8>(); run<256, 4>(); run<256, 8, 1>();
    return 0;
}
