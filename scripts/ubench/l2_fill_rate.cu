// l2_fill_rate.cu -- micro-benchmark: sustained L2 -> shared-memory fill rate of cp.async.bulk (UBLKCP) on sm_100a.
// One persistent CTA per SM keeps a ring of DEPTH copies of CHUNK bytes in flight, reading an L2-resident buffer:
//   mode 0  every CTA streams its own region (like the weight stream),
//   mode 1  all CTAs read the same small region (like the activation tile: the "hot" operand),
//   mode 2  clusters of 2: each CTA loads half a chunk and multicasts it to both CTAs of the pair,
//   mode 4  clusters of 4: each CTA loads a quarter and multicasts it to all four.
// Reports bytes landed per clock per SM (and chip-wide), from clock64 on CTA 0 and from CUDA events.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../vlut.cpp_b200/csrc/ptx.cuh"
using namespace vlut;

__device__ __forceinline__ void bulk_g2s_mc(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint16_t mask) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar), "h"(mask)
        : "memory");
}

// NW warps issue concurrently, each with its own ring of `depth` chunks (tests whether the ~277 clk per UBLKCP is per thread)
__global__ void __launch_bounds__(128, 1) fill_mw(const uint8_t *src, size_t region, int chunk, int depth, int iters, int nw,
                                                  long long *out) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full[4][16];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        for (int w = 0; w < 4; w++)
            for (int i = 0; i < depth; i++) ptx::mbar_init(ptx::smem_u32(&full[w][i]), 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0 && warp < nw) {
        const size_t base = (size_t) blockIdx.x * region;
        size_t off = (size_t) warp * chunk;
        const long long t0 = clock64();
        for (int it = 0; it < iters + depth; it++) {
            const int s = it % depth;
            const uint32_t bar = ptx::smem_u32(&full[warp][s]);
            if (it >= depth) ptx::mbar_wait(bar, ((it / depth) - 1) & 1);
            if (it < iters) {
                ptx::mbar_arrive_expect_tx(bar, chunk);
                ptx::bulk_g2s(smem_base + (warp * depth + s) * chunk, src + base + off, chunk, bar);
                off += (size_t) nw * chunk;
                if (off + chunk > region) off = (size_t) warp * chunk;
            }
        }
        const long long t1 = clock64();
        if (blockIdx.x == 0 && warp == 0) out[0] = t1 - t0;
    }
}

template <int CSZ>
__global__ void __launch_bounds__(128, 1) fill(const uint8_t *src, size_t region, int shared_src, int chunk, int depth, int iters,
                                               long long *out) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full[16];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint32_t rank = 0;
    if (CSZ > 1) rank = ptx::cluster_ctarank();
    if (threadIdx.x == 0) {
        for (int i = 0; i < depth; i++) ptx::mbar_init(ptx::smem_u32(&full[i]), 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    if (CSZ > 1) ptx::cluster_sync();
    if (threadIdx.x == 0) {
        // every CTA receives `chunk` bytes per step; with multicast it issues chunk / CSZ of them itself
        const size_t base = shared_src ? 0 : (size_t) (blockIdx.x / CSZ) * region;
        const int part = chunk / CSZ;
        const uint16_t mask = (uint16_t) ((1u << CSZ) - 1);
        size_t off = 0;
        const long long t0 = clock64();
        for (int it = 0; it < iters + depth; it++) {
            const int s = it % depth;
            const uint32_t bar = ptx::smem_u32(&full[s]);
            if (it >= depth) ptx::mbar_wait(bar, ((it / depth) - 1) & 1);  // the copy issued `depth` steps ago has landed
            if (it < iters) {
                ptx::mbar_arrive_expect_tx(bar, chunk);
                const uint8_t *p = src + base + off + (size_t) rank * part;
                if (CSZ == 1) ptx::bulk_g2s(smem_base + s * chunk, p, chunk, bar);
                else bulk_g2s_mc(smem_base + s * chunk + rank * part, p, part, bar, mask);
                off += chunk;
                if (off + chunk > region) off = 0;
            }
        }
        const long long t1 = clock64();
        if (blockIdx.x == 0) out[0] = t1 - t0;
    }
    __syncthreads();
    if (CSZ > 1) ptx::cluster_sync();
}

template <int CSZ>
static void run(const uint8_t *src, size_t region, int shared_src, int chunk, int depth, int iters, long long *dout, int sms) {
    const size_t smem = (size_t) chunk * depth + 1024;
    cudaFuncSetAttribute(fill<CSZ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    cudaLaunchConfig_t cfg{};
    int grid = sms / CSZ * CSZ;
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CSZ;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (CSZ > 1) {
        int nc = 0;
        cudaOccupancyMaxActiveClusters(&nc, fill<CSZ>, &cfg);
        if (nc * CSZ < grid) { grid = nc * CSZ; cfg.gridDim = dim3(grid); }
    }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int rep = 0; rep < 2; rep++) {  // first run warms L2
        cudaEventRecord(e0);
        cudaLaunchKernelEx(&cfg, fill<CSZ>, src, region, shared_src, chunk, depth, iters, dout);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
    }
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long clk;
    cudaMemcpy(&clk, dout, 8, cudaMemcpyDeviceToHost);
    const double bytes = (double) chunk * iters;
    printf("csz=%d src=%-7s chunk=%5d depth=%2d CTAs=%3d : %6.1f B/clk/SM (clock64, CTA 0)  chip %6.2f TB/s delivered (events)\n", CSZ,
           shared_src ? "shared" : "private", chunk, depth, grid, bytes / (double) clk, bytes * grid / (ms * 1e-3) / 1e12);
    fflush(stdout);
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t region = 512 * 1024;  // per-CTA (private) or common (shared) source window: stays L2 resident
    uint8_t *src;
    cudaMalloc(&src, region * sms);
    cudaMemset(src, 1, region * sms);
    long long *dout;
    cudaMalloc(&dout, 64);
    const int iters = 4000;
    for (int i = 0; i < 200; i++) {  // warm the clocks (~2 s)
        cudaFuncSetAttribute(fill<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        fill<1><<<sms, 128, 16384 * 8 + 1024>>>(src, region, 0, 16384, 8, iters, dout);
    }
    cudaDeviceSynchronize();
    printf("---- L2 -> shared memory fill by cp.async.bulk, %d SMs, one issuing thread per CTA ----\n", sms);
    for (int shared_src = 0; shared_src <= 1; shared_src++)
        for (int chunk : {2048, 8192, 16384, 32768, 65536})
            for (int depth : {2, 3, 6}) {
                if ((size_t) chunk * depth > 200 * 1024) continue;
                run<1>(src, region, shared_src, chunk, depth, iters, dout, sms);
            }
    printf("---- several issuing warps per CTA (each its own ring, private source) ----\n");
    cudaFuncSetAttribute(fill_mw, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int chunk : {2048, 8192, 16384})
        for (int nw : {1, 2, 4}) {
            const int depth = 3;
            if ((size_t) chunk * depth * nw > 200 * 1024) continue;
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0);
            cudaEventCreate(&e1);
            for (int rep = 0; rep < 2; rep++) {
                cudaEventRecord(e0);
                fill_mw<<<sms, 128, (size_t) chunk * depth * nw + 1024>>>(src, region, chunk, depth, iters, nw, dout);
                cudaEventRecord(e1);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            }
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            long long clk;
            cudaMemcpy(&clk, dout, 8, cudaMemcpyDeviceToHost);
            printf("warps=%d chunk=%5d depth=%d : %6.1f B/clk/SM (clock64, CTA 0)  chip %6.2f TB/s (events)\n", nw, chunk, depth,
                   (double) chunk * iters * nw / (double) clk, (double) chunk * iters * nw * sms / (ms * 1e-3) / 1e12);
            fflush(stdout);
        }
    return 0;
}
