// stream_once.cu -- how fast can ONE kernel launch pull a 14.7 MB weight matrix (4096 -> 14336, 2 bits per weight) out of HBM
// into the SMs, whatever it then does with it?  The floor under the small-T GEMM (gemm_tcs.cu), whose timeline shows the last
// weight bytes landing ~6000 clk after the kernel starts.
//   mode 0  per-thread cp.async 16 B, 512 threads per CTA, everything requested at once (the TCS pattern), CTAs = strips
//   mode 1  cp.async.bulk, one thread per CTA, chunks of CH bytes, all requested at once
//   mode 2  ld.global.v4 (LDG.128), 512 threads, UNR loads in flight per thread
// for grids of 112 and 148 (and 296) CTAs.  12 buffers are rotated so that nothing is an L2 hit (176 MB > 126 MB L2).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_once stream_once.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../vlut.cpp_b200/csrc/ptx.cuh"
using namespace vlut;

constexpr size_t BYTES = (size_t) 4096 * 14336 / 4;   // 14,680,064
constexpr int NBUF = 12;

__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(const uint8_t *__restrict__ src, size_t per_cta, int chunk, unsigned long long *tim,
                                            uint32_t *sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    const unsigned long long t0 = gtime();
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");   // (only matters for launches with the PDL attribute)
    const uint8_t *base = src + (size_t) blockIdx.x * per_cta;
    const uint32_t sbase = ptx::smem_u32(smem);
    uint32_t acc = 0;
    if (MODE == 0) {
        // 512 threads x 16 B = 8 KB per step
        for (size_t o = (size_t) threadIdx.x * 16; o < per_cta; o += 512 * 16) ptx::cp_async16(sbase + (uint32_t) o, base + o);
        ptx::cp_async_commit();
        ptx::cp_async_wait<0>();
        __syncthreads();
    } else if (MODE == 1) {
        if (threadIdx.x == 0) {
            ptx::mbar_init(ptx::smem_u32(&bar), 1);
            ptx::fence_barrier_init();
            ptx::mbar_arrive_expect_tx(ptx::smem_u32(&bar), (uint32_t) per_cta);
            for (size_t o = 0; o < per_cta; o += chunk)
                ptx::bulk_g2s(sbase + (uint32_t) o, base + o, (uint32_t) min((size_t) chunk, per_cta - o), ptx::smem_u32(&bar));
        }
        __syncthreads();
        ptx::mbar_wait(ptx::smem_u32(&bar), 0);
    } else {
        const uint4 *p4 = reinterpret_cast<const uint4 *>(base);
        const size_t n4 = per_cta / 16;
#pragma unroll 8
        for (size_t i = threadIdx.x; i < n4; i += 512) {
            const uint4 v = __ldcs(p4 + i);
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    }
    if (MODE != 2) acc = ((volatile uint32_t *) smem)[threadIdx.x];
    if (acc == 0x12345678u) sink[0] = acc;
    __syncthreads();
    if (threadIdx.x == 0) { tim[2 * blockIdx.x] = t0; tim[2 * blockIdx.x + 1] = gtime(); }
}

template <int MODE>
void run(const char *name, int grid, int chunk, uint8_t *bufs, unsigned long long *tim, uint32_t *sink, bool pdl = false) {
    size_t per_cta = (BYTES / grid) & ~(size_t) 1023;   // a multiple of 1 KB (the few bytes left over do not matter here)
    if (per_cta > 200 * 1024) { printf("%s grid %d: %zu B per CTA does not fit shared memory\n", name, grid, per_cta); return; }
    auto kern = k<MODE>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const size_t sm = MODE == 2 ? 0 : per_cta;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto launch = [&](int i) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(512); cfg.dynamicSmemBytes = sm;
        cudaLaunchAttribute at{};
        at.id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at.val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = &at; cfg.numAttrs = pdl ? 1 : 0;
        cudaLaunchKernelEx(&cfg, kern, (const uint8_t *) (bufs + i * BYTES), per_cta, chunk, tim, sink);
    };
    for (int i = 0; i < NBUF; i++) launch(i);
    cudaDeviceSynchronize();
    float best = 1e9f, sum = 0;
    double span_sum = 0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        for (int i = 0; i < NBUF; i++) launch(i);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        best = ms < best ? ms : best; sum += ms;
        static unsigned long long h[2 * 512];
        cudaMemcpy(h, tim, 2 * grid * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        unsigned long long a = ~0ull, b = 0;
        for (int i = 0; i < grid; i++) { a = h[2 * i] < a ? h[2 * i] : a; b = h[2 * i + 1] > b ? h[2 * i + 1] : b; }
        span_sum += (double) (b - a);
    }
    const double us = best / NBUF * 1e3, span = span_sum / 5 / 1e3;
    printf("%s%-22s grid %3d chunk %6d: %6.2f us per launch back to back (%5.0f GB/s), first CTA start -> last CTA end %5.2f us (%5.0f GB/s)\n",
           pdl ? "PDL " : "    ", name, grid, chunk, us, grid * per_cta / us / 1e3, span, grid * per_cta / span / 1e3);
}

int main() {
    uint8_t *bufs; unsigned long long *tim; uint32_t *sink;
    cudaMalloc(&bufs, NBUF * BYTES); cudaMemset(bufs, 1, NBUF * BYTES);
    cudaMalloc(&tim, 2 * 512 * sizeof(unsigned long long)); cudaMalloc(&sink, 4);
    for (int grid : {112, 148}) {
        run<0>("cp.async 16B/thread", grid, 0, bufs, tim, sink);
        for (int ch : {2048, 8192, 32768}) run<1>("cp.async.bulk", grid, ch, bufs, tim, sink);
        run<2>("ld.global.v4 x8", grid, 0, bufs, tim, sink);
    }
    run<2>("ld.global.v4 x8", 296, 0, bufs, tim, sink);
    // the same launches with programmatic dependent launch (the next kernel may start while this one runs -- on SMs that have
    // room for its CTA: none with 128 KB of shared memory per CTA beyond the 36 SMs a 112-CTA grid leaves idle)
    run<0>("cp.async 16B/thread", 112, 0, bufs, tim, sink, true);
    run<0>("cp.async 16B/thread", 148, 0, bufs, tim, sink, true);
    run<1>("cp.async.bulk", 112, 8192, bufs, tim, sink, true);
    run<2>("ld.global.v4 x8", 112, 0, bufs, tim, sink, true);
    run<2>("ld.global.v4 x8", 296, 0, bufs, tim, sink, true);
    return 0;
}
