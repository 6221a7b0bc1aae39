// pair_pingpong.cu -- hand-off latency of the CTA-pair kernel's A ring, in isolation (cluster of 2).
//   "decode" warps (4 per CTA): wait aempty[i % NA] (local mbarrier) -> lane 0 arrives on the LEADER's afull[i % NA]
//   "issuer" warp (leader CTA): wait afull[i % NA] (8 arrivals) -> tcgen05.commit ... multicast (or two plain arrives) on
//   aempty[i % NA] of both CTAs.
// With NA = 1 the period is the bare round trip decode -> issuer -> decode; NA = 2, 4 show how much of it pipelines.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pair_pingpong pair_pingpong.cu ; run: ./pair_pingpong
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ uint32_t mapa(uint32_t a, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
    return r;
}
__device__ __forceinline__ void arrive_cluster(uint32_t a) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(a) : "memory");
}
__device__ __forceinline__ void arrive_local(uint32_t a) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(a) : "memory");
}
__device__ __forceinline__ void commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((uint16_t) 3) : "memory");
}
__device__ __forceinline__ uint32_t ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// MODE 0: tcgen05.commit multicast; 1: two plain arrives (local + remote); 2: as 0 but only the LEADER's decode warps take
// part (afull count 4: no remote arrive on the way in)
template <int NA, int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1) pingpong(int iters, long long *out) {
    __shared__ __align__(8) uint64_t afull[4], aempty[4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = ctarank();
    if (threadIdx.x == 0) {
        for (int i = 0; i < NA; i++) {
            mbar_init(smem_u32(&afull[i]), MODE == 2 ? 4 : 8);
            mbar_init(smem_u32(&aempty[i]), 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    cluster_sync();
    const long long t0 = clock64();
    if (warp < 4) {   // decode warps
        if (MODE != 2 || rank == 0) {
            const uint32_t afull_leader = mapa(smem_u32(&afull[0]), 0);
            int i = 0;
            uint32_t ph = 0;
            for (int it = 0; it < iters; it++) {
                mbar_wait(smem_u32(&aempty[i]), ph ^ 1);
                __syncwarp();
                if (lane == 0) arrive_cluster(afull_leader + 8u * i);
                if (++i == NA) { i = 0; ph ^= 1; }
            }
        }
    } else if (warp == 4 && rank == 0) {   // issuer
        int i = 0;
        uint32_t ph = 0;
        for (int it = 0; it < iters; it++) {
            mbar_wait(smem_u32(&afull[i]), ph);
            if (lane == 0) {
                if (MODE == 1) {
                    arrive_local(smem_u32(&aempty[i]));
                    arrive_cluster(mapa(smem_u32(&aempty[i]), 1));
                } else {
                    commit_pair(smem_u32(&aempty[i]));
                }
            }
            __syncwarp();
            if (++i == NA) { i = 0; ph ^= 1; }
        }
    }
    __syncthreads();
    cluster_sync();
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = clock64() - t0;
}

template <int NA, int MODE>
static void run(const char *name) {
    long long *d;
    cudaMalloc(&d, 8);
    const int iters = 20000;
    pingpong<NA, MODE><<<2, 192>>>(iters, d);
    pingpong<NA, MODE><<<2, 192>>>(iters, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h = 0;
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("%-58s NA=%d: %7.1f clk per hand-off  (%s)\n", name, NA, (double) h / iters, cudaGetErrorString(e));
    cudaFree(d);
}

int main() {
    run<1, 0>("decode(both CTAs) -> issuer -> tcgen05.commit multicast");
    run<2, 0>("decode(both CTAs) -> issuer -> tcgen05.commit multicast");
    run<4, 0>("decode(both CTAs) -> issuer -> tcgen05.commit multicast");
    run<1, 1>("decode(both CTAs) -> issuer -> plain arrives");
    run<2, 1>("decode(both CTAs) -> issuer -> plain arrives");
    run<4, 1>("decode(both CTAs) -> issuer -> plain arrives");
    run<1, 2>("decode(leader only) -> issuer -> tcgen05.commit multicast");
    run<4, 2>("decode(leader only) -> issuer -> tcgen05.commit multicast");
    return 0;
}
