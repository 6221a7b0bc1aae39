// tmem_st_rate.cu -- micro-benchmark: tcgen05.st (registers -> tensor memory) throughput on sm_100a, and the rate of
// the full decode step of the Vec-LUT kernel (LDS.128 of packed 2-bit codes -> shift/mask -> tcgen05.st) without MMA.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../vlut.cpp_b200/csrc/ptx.cuh"
using namespace vlut;

template <int NW, int MODE>  // MODE 0: st.x16 only; 1: st.x32; 2: LDS + decode + st.x16 (the kernel's decode step)
__global__ void __launch_bounds__(NW * 32, 1) k(long long *out, int iters) {
    __shared__ __align__(16) uint4 packed[128 * 8];
    __shared__ uint32_t tbase;
    for (int i = threadIdx.x; i < 128 * 8; i += NW * 32) packed[i] = make_uint4(i * 2654435761u, i * 40503u, ~i, i ^ 0x55aa55aau);
    if (threadIdx.x < 32) { ptx::tmem_alloc(ptx::smem_u32(&tbase), 512); ptx::tmem_relinquish(); }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t taddr = tbase + ((uint32_t) ((warp & 3) * 32) << 16) + (warp >> 2) * 128;
    uint32_t r[32];
    for (int i = 0; i < 32; i++) r[i] = threadIdx.x * 31 + i;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) {
#pragma unroll
            for (int c = 0; c < 8; c++) {
                uint32_t v[16];
#pragma unroll
                for (int i = 0; i < 16; i++) v[i] = r[i] + c + it;
                ptx::tmem_st_32x32b_x16(taddr + c * 16, v);
            }
        } else if (MODE == 1) {
#pragma unroll
            for (int c = 0; c < 4; c++) {
                asm volatile(
                    "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
                    "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr + c * 32),
                    "r"(r[0] + it), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
                    "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
                    "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
                    "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31] + c)
                    : "memory");
            }
        } else {
            const uint32_t base = ptx::smem_u32(packed) + ((warp & 3) * 32 + lane) * 16;
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const uint4 w = ptx::lds128(base + c * 2048);
                const uint32_t wv[4] = {w.x, w.y, w.z, w.w};
                uint32_t v[16];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    v[4 * j + 0] = wv[j] & 0x03030303u;
                    v[4 * j + 1] = __umulhi(wv[j], 1u << 30) & 0x03030303u;
                    v[4 * j + 2] = __umulhi(wv[j], 1u << 28) & 0x03030303u;
                    v[4 * j + 3] = __umulhi(wv[j], 1u << 26) & 0x03030303u;
                }
                ptx::tmem_st_32x32b_x16(taddr + c * 16, v);
            }
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    ptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tbase, 512); }
}

template <int NW, int MODE>
void run(const char *name) {
    long long *d; cudaMalloc(&d, 148 * sizeof(long long));
    const int iters = 200;
    k<NW, MODE><<<148, NW * 32>>>(d, iters);
    k<NW, MODE><<<148, NW * 32>>>(d, iters);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double cyc = 0; for (int i = 0; i < 148; i++) cyc += h[i]; cyc /= 148;
    const double bytes = (double) NW * 32 * 128 * 4;  // per iteration: every thread writes 128 columns x 4 B
    printf("%-28s warps=%d: %.0f cyc/iter -> %.1f B/clk/SM TMEM write (%s)\n", name, NW, cyc / iters, bytes / (cyc / iters),
           cudaGetErrorString(e));
    cudaFree(d);
}


// decode step (as MODE 2) while one more warp keeps the tensor pipe busy with TS-mode MMAs reading the other TMEM columns
__device__ __forceinline__ uint64_t mk_desc(uint32_t a) {
    uint64_t d = 0;
    d |= (uint64_t) ((a & 0x3FFFFu) >> 4);
    d |= (uint64_t) 1 << 16;
    d |= (uint64_t) (1024 >> 4) << 32;
    d |= (uint64_t) 1 << 46;
    d |= (uint64_t) 2 << 61;
    return d;
}
template <int NW, int N, int NSPIN, int WAITMODE>
__global__ void __launch_bounds__(NW * 32 + 32 + NSPIN * 32, 1) kmix(long long *out, int iters, int with_mma) {
    extern __shared__ __align__(1024) uint8_t dyn[];
    __shared__ __align__(16) uint4 packed[128 * 8];
    __shared__ uint32_t tbase;
    __shared__ __align__(8) uint64_t bar, spinbar;
    __shared__ volatile int stop;
    const uint32_t bsm = (ptx::smem_u32(dyn) + 1023u) & ~1023u;
    for (int i = threadIdx.x; i < 128 * 8; i += blockDim.x) packed[i] = make_uint4(i * 2654435761u, i * 40503u, ~i, i ^ 0x55aa55aau);
    for (int i = threadIdx.x; i < 16384 / 4; i += blockDim.x) ((uint32_t *) dyn)[i + 256] = 0x01010101u;
    if (threadIdx.x == 0) { stop = 0; ptx::mbar_init(ptx::smem_u32(&bar), 1); ptx::mbar_init(ptx::smem_u32(&spinbar), 1); ptx::fence_barrier_init(); }
    if (threadIdx.x < 32) { ptx::tmem_alloc(ptx::smem_u32(&tbase), 512); ptx::tmem_relinquish(); }
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp > NW) {
        // waits until the decode warps are done: 0 plain try_wait loop, 1 try_wait with a suspend-time hint, 2 nanosleep back-off
        const uint32_t sb = ptx::smem_u32(&spinbar);
        if (WAITMODE == 0) {
            ptx::mbar_wait(sb, 0);
        } else if (WAITMODE == 1) {
            uint32_t ok;
            do {
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(ok) : "r"(sb), "r"(0u), "r"(0x989680u) : "memory");
            } while (!ok);
        } else {
            while (!ptx::mbar_try_wait(sb, 0)) __nanosleep(64);
        }
    } else if (warp == NW) {
        if (with_mma) {
            constexpr uint32_t idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t) (N >> 3) << 17) | ((uint32_t) (128 >> 4) << 24);
            const uint64_t b_desc = mk_desc(bsm);
            uint32_t ph = 0;
            while (!stop) {
                if (ptx::elect_one()) {
                    for (int j = 0; j < 16; j++) ptx::umma_i8_ts(tbase + 256, tbase + 384 + (j & 3) * 8, b_desc + (j & 3) * 2, idesc, 1);
                    ptx::umma_commit(ptx::smem_u32(&bar));
                }
                __syncwarp();
                ptx::mbar_wait(ptx::smem_u32(&bar), ph);
                ph ^= 1;
            }
        }
    } else {
        const uint32_t taddr = tbase + ((uint32_t) ((warp & 3) * 32) << 16) + (warp >> 2) * 128;
        const uint32_t base = ptx::smem_u32(packed) + ((warp & 3) * 32 + lane) * 16;
        const long long t0 = clock64();
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int c = 0; c < 8; c++) {
                uint4 w = ptx::lds128(base + c * 2048);
                w.x += it;  // keep the unpack inside the loop
                const uint32_t wv[4] = {w.x, w.y, w.z, w.w};
                uint32_t v[16];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    v[4 * j + 0] = wv[j] & 0x03030303u;
                    v[4 * j + 1] = __umulhi(wv[j], 1u << 30) & 0x03030303u;
                    v[4 * j + 2] = __umulhi(wv[j], 1u << 28) & 0x03030303u;
                    v[4 * j + 3] = __umulhi(wv[j], 1u << 26) & 0x03030303u;
                }
                ptx::tmem_st_32x32b_x16(taddr + c * 16, v);
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        }
        const long long t1 = clock64();
        if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
        asm volatile("bar.sync 1, %0;" ::"r"(NW * 32));
        if (threadIdx.x == 0) { stop = 1; ptx::mbar_arrive(ptx::smem_u32(&spinbar)); }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tbase, 512); }
}
template <int NW, int N, int NSPIN, int WAITMODE = 0>
void runmix(int with_mma) {
    long long *d; cudaMalloc(&d, 148 * sizeof(long long));
    const int iters = 200;
    cudaFuncSetAttribute(kmix<NW, N, NSPIN, WAITMODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 1024);
    for (int r = 0; r < 2; r++) kmix<NW, N, NSPIN, WAITMODE><<<148, NW * 32 + 32 + NSPIN * 32, 40 * 1024>>>(d, iters, with_mma);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double cyc = 0; for (int i = 0; i < 148; i++) cyc += h[i]; cyc /= 148;
    printf("decode step, %d warps x 8 chunks, %2d waiting warps (mode %d), concurrent TS MMA N=%3d %s: %.0f clk per iteration (%s)\n", NW, NSPIN, WAITMODE, N, with_mma ? "ON " : "off",
           cyc / iters, cudaGetErrorString(e));
    cudaFree(d);
}

int main() {
    run<4, 0>("st.32x32b.x16");
    run<8, 0>("st.32x32b.x16");
    run<4, 1>("st.32x32b.x32");
    run<8, 1>("st.32x32b.x32");
    run<4, 2>("LDS+decode+st.x16");
    run<8, 2>("LDS+decode+st.x16");
    runmix<4, 16, 0>(0); runmix<4, 16, 0>(1); runmix<8, 16, 0>(0); runmix<8, 16, 0>(1);
    runmix<4, 128, 0>(1); runmix<8, 128, 0>(1);
    runmix<4, 16, 4>(1); runmix<4, 16, 12>(1); runmix<8, 16, 8>(1);
    runmix<4, 16, 12, 1>(1); runmix<8, 16, 8, 1>(1); runmix<4, 16, 12, 2>(1); runmix<8, 16, 8, 2>(1);
    return 0;
}
