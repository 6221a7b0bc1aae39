// peer_store.cu -- how fast can ONE GPU push a [T][N_local] fp32 slice into a PEER's [T][N_total] buffer over NVLink, per
// store pattern?  (The fused all-gather epilogue of the GEMM is exactly this traffic.)   One process, two devices.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o peer_store peer_store.cu && ./peer_store
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int T = 512, NL = 14336, NTOT = 2 * NL;   // the slice: T rows of NL floats, row stride NTOT floats at the destination

// pattern A: the epilogue's: a warp instruction = 4 tokens x 128 contiguous bytes (lane = 16 B), tiles of 128 rows x 16 tokens
__global__ void k_epi(float *dst, int iters, int threads_active) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
    // work items: (row tile of 128 floats = 4 warps-worth of 32 floats, token group of 4)
    const int row_tiles = NL / 32, tok_groups = T / 4;
    const long total = (long) row_tiles * tok_groups;
    for (int it = 0; it < iters; it++)
        for (long w = (long) blockIdx.x * nw + warp; w < total; w += (long) gridDim.x * nw) {
            const int rt = (int) (w % row_tiles), tg = (int) (w / row_tiles);
            const int tok = tg * 4 + (lane & 3) ;
            const int col = rt * 32 + (lane >> 2) * 4;
            *reinterpret_cast<float4 *>(dst + (long) tok * NTOT + col) = v;
        }
}
// pattern B: a warp instruction = 512 contiguous bytes of one token
__global__ void k_row(float *dst, int iters) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
    const int segs = NL / 128;
    const long total = (long) segs * T;
    for (int it = 0; it < iters; it++)
        for (long w = (long) blockIdx.x * nw + warp; w < total; w += (long) gridDim.x * nw) {
            const int sg = (int) (w % segs), tok = (int) (w / segs);
            *reinterpret_cast<float4 *>(dst + (long) tok * NTOT + sg * 128 + lane * 4) = v;
        }
}
// pattern C: bulk async copies shared -> peer global, `bytes` per copy (one row segment), issued by every lane of warp 0..nw-1
__global__ void k_bulk(float *dst, int iters, int bytes) {
    extern __shared__ __align__(128) uint8_t smem[];
    for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) reinterpret_cast<float *>(smem)[i] = 1.f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const uint32_t src = (uint32_t) __cvta_generic_to_shared(smem);
    const int fl = bytes / 4, segs = NL / fl;
    const long total = (long) segs * T;
    for (int it = 0; it < iters; it++) {
        for (long w = (long) blockIdx.x * blockDim.x + threadIdx.x; w < total; w += (long) gridDim.x * blockDim.x) {
            const int sg = (int) (w % segs), tok = (int) (w / segs);
            float *d = dst + (long) tok * NTOT + sg * fl;
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(d), "r"(src), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 8;" ::: "memory");
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <typename F>
static float timeit(F f, cudaStream_t s) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f(); CK(cudaStreamSynchronize(s));
    CK(cudaEventRecord(e0, s));
    f();
    CK(cudaEventRecord(e1, s));
    CK(cudaStreamSynchronize(s));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    return ms;
}

int main() {
    int nd = 0; CK(cudaGetDeviceCount(&nd));
    float *local, *peer;
    CK(cudaSetDevice(0));
    CK(cudaMalloc(&local, (size_t) T * NTOT * 4));
    if (nd > 1) {
        CK(cudaSetDevice(1)); CK(cudaMalloc(&peer, (size_t) T * NTOT * 4));
        CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0));
    } else peer = local;
    cudaStream_t s; CK(cudaStreamCreate(&s));
    const int iters = 8;
    const double mb = (double) T * NL * 4 * iters / 1e6;
    for (int where = 0; where < 2; where++) {
        float *d = where ? peer : local;
        const char *nm = where ? (nd > 1 ? "peer " : "local") : "local";
        for (int thr : {128, 256, 512})
            for (int grid : {148, 296}) {
                float a = timeit([&] { k_epi<<<grid, thr, 0, s>>>(d, iters, thr); }, s);
                float b = timeit([&] { k_row<<<grid, thr, 0, s>>>(d, iters); }, s);
                printf("%s grid %3d x %3d thr: 4x128B-per-instr %7.1f GB/s   512B-per-instr %7.1f GB/s\n", nm, grid, thr, mb / a, mb / b);
            }
        for (int bytes : {512, 2048, 8192})
            for (int thr : {32, 128}) {
                CK(cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
                float c = timeit([&] { k_bulk<<<148, thr, bytes, s>>>(d, iters, bytes); }, s);
                printf("%s bulk s2g %5d B x %3d issuing threads/CTA: %7.1f GB/s\n", nm, bytes, thr, mb / c);
            }
        float m = timeit([&] { for (int i = 0; i < iters; i++) CK(cudaMemcpy2DAsync(d, (size_t) NTOT * 4, local + NL, (size_t) NTOT * 4, (size_t) NL * 4, T, cudaMemcpyDeviceToDevice, s)); }, s);
        printf("%s cudaMemcpy2DAsync: %7.1f GB/s\n", nm, mb / m);
    }
    return 0;
}
