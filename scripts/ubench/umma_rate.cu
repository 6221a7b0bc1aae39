// umma_rate.cu -- micro-benchmark: issue rate / throughput of tcgen05.mma kind::i8 (M=128) on sm_100a
// for several N, with the A operand in TMEM (TS) or shared memory (SS). One CTA per SM, one issuing thread.
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
__host__ __device__ constexpr uint32_t make_idesc(int n, int fmt_f16) {
    return fmt_f16 ? ((1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (128 >> 4) << 24))
                   : ((2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t) (n >> 3) << 17) | ((uint32_t) (128 >> 4) << 24));
}
__device__ __forceinline__ void umma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                 "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}

template <int N, int MODE>  // MODE 0: i8 SS, 1: i8 TS, 2: bf16 SS
__global__ void __launch_bounds__(128, 1) k(long long *out, int iters, int nacc) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tbase;
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 128) ((uint32_t *) smem_raw)[i] = 0x01010101u;
    if (threadIdx.x == 0) { ptx::mbar_init(ptx::smem_u32(&bar), 1); ptx::fence_barrier_init(); }
    if (threadIdx.x < 32) { ptx::tmem_alloc(ptx::smem_u32(&tbase), 512); ptx::tmem_relinquish(); }
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = tbase;
    if (threadIdx.x < 32) {
        const uint64_t a_desc = make_smem_desc(smem_base);
        const uint64_t b_desc = make_smem_desc(smem_base + 16384);
        constexpr uint32_t idesc = make_idesc(N, MODE == 2);
        long long t0 = 0, t1 = 0, t2 = 0;
        if (ptx::elect_one()) {
            t0 = clock64();
            for (int it = 0; it < iters; it++) {
                const uint32_t d = tm + (it % nacc) * N;
#pragma unroll
                for (int ks = 0; ks < 4; ks++) {
                    if (MODE == 1) ptx::umma_i8_ts(d, tm + 384 + ks * 8, b_desc + ks * 2, idesc, 1);
                    else if (MODE == 0) ptx::umma_i8_ss(d, a_desc + ks * 2, b_desc + ks * 2, idesc, 1);
                    else umma_f16_ss(d, a_desc + ks * 2, b_desc + ks * 2, idesc, 1);
                }
            }
            t1 = clock64();
            ptx::umma_commit(ptx::smem_u32(&bar));
        }
        __syncwarp();
        ptx::mbar_wait(ptx::smem_u32(&bar), 0);
        t2 = clock64();
        if (threadIdx.x == 0) { out[blockIdx.x * 2] = t1 - t0; out[blockIdx.x * 2 + 1] = t2 - t0; }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tm, 512); }
}

template <int N, int MODE>
void run(const char *name, int nacc) {
    long long *d; cudaMalloc(&d, 148 * 2 * sizeof(long long));
    const int iters = 256;
    auto kern = k<N, MODE>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int grid : {1, 148}) {
        kern<<<grid, 128, 64 * 1024>>>(d, iters, nacc);
        kern<<<grid, 128, 64 * 1024>>>(d, iters, nacc);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[296]; cudaMemcpy(h, d, grid * 2 * sizeof(long long), cudaMemcpyDeviceToHost);
        double issue = 0, total = 0; for (int i = 0; i < grid; i++) { issue += h[2 * i]; total += h[2 * i + 1]; }
        issue /= grid; total /= grid;
        const double macs = 128.0 * N * (MODE == 2 ? 16 : 32);
        printf("%-10s N=%3d nacc=%d grid=%3d: issue %.1f cyc/MMA, complete %.1f cyc/MMA -> %.0f MAC/cyc/SM  (%s)\n", name, N, nacc, grid,
               issue / (iters * 4), total / (iters * 4), macs / (total / (iters * 4)), cudaGetErrorString(e));
    }
    cudaFree(d);
}

int main() {
    run<16, 0>("i8 SS", 1);  run<16, 1>("i8 TS", 1);  run<16, 1>("i8 TS", 2);
    run<64, 0>("i8 SS", 1);  run<64, 1>("i8 TS", 1);
    run<128, 0>("i8 SS", 1); run<128, 1>("i8 TS", 1); run<128, 1>("i8 TS", 2);
    run<256, 0>("i8 SS", 1); run<256, 1>("i8 TS", 1);
    run<128, 2>("bf16 SS", 1); run<256, 2>("bf16 SS", 1);
    return 0;
}
