// barrier.cu -- the cross-GPU barrier that publishes a fused all-gather (vlut_mul_mat_scatter) -- the repo's own kernel.
//
// Every rank owns a small flag array in peer-visible memory (a symmetric-memory signal pad, or a peer-mapped allocation of
// a vlut_group).  Arrival = one st.release.sys of the call's epoch into slot [my rank] of EVERY rank's array; departure =
// every slot of my own array has reached the epoch.  The scatter GEMM that precedes it on the stream has completed when
// this kernel starts (plain launch: full stream serialisation); the system-scope fence in front of the flag stores orders
// its NVLink stores (peer st / multimem.st) before the flags.  The epoch lives in word [8] of the rank's own array and is
// advanced by the kernel itself, so the barrier can be captured in a CUDA graph and replayed (every rank runs the same
// number of barriers); epochs only grow, so the flags never need clearing.
#include "vlut_common.cuh"

namespace vlut {

struct BarrierArgs {
    uint32_t *flags[8];   // flags[i] = rank i's flag array (>= 8 words), as seen from this device
    int n, rank;
};

__global__ void __launch_bounds__(32) barrier_kernel(BarrierArgs a) {
    const int i = threadIdx.x;
    uint32_t epoch = 0;
    if (i == 0) {
        epoch = a.flags[a.rank][8] + 1;
        a.flags[a.rank][8] = epoch;
    }
    epoch = __shfl_sync(0xffffffffu, epoch, 0);
    if (i >= a.n) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.flags[i] + a.rank), "r"(epoch) : "memory");
    uint32_t v;
    do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.flags[a.rank] + i) : "memory");
    } while ((int32_t) (v - epoch) < 0);
}

cudaError_t launch_barrier(uint32_t *const *flags, int n, int rank, cudaStream_t s) {
    if (n < 1 || n > 8 || rank < 0 || rank >= n) return cudaErrorInvalidValue;
    BarrierArgs a{};
    for (int i = 0; i < n; i++) a.flags[i] = flags[i];
    a.n = n;
    a.rank = rank;
    barrier_kernel<<<1, 32, 0, s>>>(a);
    return cudaGetLastError();
}

}  // namespace vlut
