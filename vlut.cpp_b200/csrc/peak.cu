// peak.cu -- vlut_measure_int8_peak: the int8 tensor-pipe peak of THIS device, measured, not assumed.
//
// One CTA per SM, one elected thread issuing tcgen05.mma kind::i8 (M = 128, N = 256, K = 32, A operand in TMEM as in the
// GEMM kernels) back to back on operands that never change; the whole grid is timed with CUDA events on the context's
// stream.  This is the denominator of bench.py's `frac_int8` ("measure with cuBLASLt int8 or own
// tcgen05 micro-bench").  On B200 at 1965 MHz the loop sustains 127.5 clk per MMA = 8158 MAC/clk/SM (profiles/).
#include <vector>

#include "ptx.cuh"
#include "vlut_common.cuh"

namespace vlut {
namespace peak {

constexpr int N = 256;

__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {  // K-major, SWIZZLE_128B, 8-row atoms 1024 B apart
    uint64_t d = 0;
    d |= (uint64_t) ((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t) 1 << 16;
    d |= (uint64_t) (1024 >> 4) << 32;
    d |= (uint64_t) 1 << 46;
    d |= (uint64_t) 2 << 61;
    return d;
}

__global__ void __launch_bounds__(128, 1) kernel(int iters, long long *clocks) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tbase;
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    for (int i = threadIdx.x; i < (N * 128) / 4; i += 128) reinterpret_cast<uint32_t *>(smem_raw + (smem_base - ptx::smem_u32(smem_raw)))[i] = 0x01010101u;
    if (threadIdx.x == 0) {
        ptx::mbar_init(ptx::smem_u32(&bar), 1);
        ptx::fence_barrier_init();
    }
    if (threadIdx.x < 32) {
        ptx::tmem_alloc(ptx::smem_u32(&tbase), 512);
        ptx::tmem_relinquish();
    }
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = tbase;
    if (threadIdx.x < 32) {
        const uint64_t b_desc = make_smem_desc(smem_base);
        // D s32, A u8 (TMEM), B s8, K-major, M = 128, N = 256
        constexpr uint32_t idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t) (N >> 3) << 17) | ((uint32_t) (128 >> 4) << 24);
        long long t0 = 0;
        if (ptx::elect_one()) {
            t0 = clock64();
            for (int it = 0; it < iters; it++) {
#pragma unroll
                for (int ks = 0; ks < 4; ks++) ptx::umma_i8_ts(tm, tm + 384 + ks * 8, b_desc + ks * 2, idesc, 1);
            }
            ptx::umma_commit(ptx::smem_u32(&bar));
        }
        __syncwarp();
        ptx::mbar_wait(ptx::smem_u32(&bar), 0);
        if (threadIdx.x == 0 && clocks) clocks[blockIdx.x] = clock64() - t0;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tm, 512);
    }
}

}  // namespace peak

// TOP/s (2 * MAC) over the whole device and the average SM clocks per MMA; `iters` x 4 MMAs per SM
cudaError_t measure_int8_peak(int sm_count, cudaStream_t s, int iters, double *tops, double *clk_per_mma) {
    const size_t smem = peak::N * 128 + 1024;
    cudaError_t e = cudaFuncSetAttribute(peak::kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    if (e != cudaSuccess) return e;
    long long *clocks = nullptr;
    if ((e = cudaMalloc(&clocks, sm_count * sizeof(long long))) != cudaSuccess) return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best_ms = 1e30;
    for (int rep = 0; rep < 4; rep++) {   // first repetition = warm-up
        cudaEventRecord(e0, s);
        peak::kernel<<<sm_count, 128, smem, s>>>(iters, clocks);
        cudaEventRecord(e1, s);
        if ((e = cudaEventSynchronize(e1)) != cudaSuccess) break;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best_ms) best_ms = ms;
    }
    if (e == cudaSuccess) {
        std::vector<long long> h(sm_count);
        e = cudaMemcpy(h.data(), clocks, sm_count * sizeof(long long), cudaMemcpyDeviceToHost);
        double c = 0;
        for (long long v : h) c += (double) v;
        if (clk_per_mma) *clk_per_mma = c / sm_count / (4.0 * iters);
        if (tops) *tops = 2.0 * 128 * peak::N * 32 * 4.0 * iters * sm_count / (best_ms * 1e-3) / 1e12;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(clocks);
    return e;
}

}  // namespace vlut
