// repack.cu -- GGUF I1/I2 whole-matrix packing  ->  device "T2" tiled 2-bit layout (see vlut_common.cuh).
//
// Source layouts (reference ggml/src/ggml-quants-vlut.c): byte p of row n holds 4 (I2*) or 5 (I1*, first
// nb5 = K/20*4 bytes; then (K%20)/4 four-trit tail bytes) base-3 digits d_j = w+1, little-endian; byte p of
// row n lives at  (p/il)*il*N + n*il + p%il  with interleave il = 1 (I2_V, I1_V), 4 (I2_V_4), 8 (I2_V_8),
// 2 (I1_V_2, five-trit bytes only; its tail bytes are plain p*N + n).
//
// One thread produces the 16 bytes (64 consecutive k) of one row of one tile.  HBM-bound byte shuffle: reads
// are gathers out of the N-fastest GGUF layout (consecutive rows -> consecutive bytes, so consecutive threads
// hit consecutive addresses), writes are fully coalesced 128-bit stores.  Runs once per tensor at load.
#include "vlut_common.cuh"

namespace vlut {

__device__ __forceinline__ int64_t gguf_byte_pos(int il, int64_t p, int64_t n, int64_t N) {
    return (p / il) * il * N + n * il + p % il;
}

// decode code (w+1) of element k of row n from the GGUF buffer
template <bool I1>
__device__ __forceinline__ uint32_t gguf_code(const uint8_t *__restrict__ src, int il, int64_t n, int64_t k, int64_t N,
                                              int64_t nb5) {
    int64_t p;
    int d;
    int ilp = il;
    if (I1) {
        if (k < nb5 * 5) { p = k / 5; d = (int) (k % 5); }
        else { p = nb5 + (k - nb5 * 5) / 4; d = (int) ((k - nb5 * 5) % 4); ilp = 1; }
    } else { p = k / 4; d = (int) (k % 4); }
    uint32_t b = src[gguf_byte_pos(ilp, p, n, N)];
    // digit d of b in base 3
    const uint32_t pw[5] = {1, 3, 9, 27, 81};
    return (b / pw[d]) % 3u;
}

// grid: one thread per 16-byte row entry of a (strip, 64-k chunk) tile; a warp writes 512 contiguous bytes
template <bool I1>
__global__ void __launch_bounds__(256) repack_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, int il,
                                                     int64_t K, int64_t N, int64_t Kpad, int64_t Npad) {
    const int64_t gid = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = Npad * Kpad / 4 / 16;
    if (gid >= total) return;
    const int r = (int) (gid % TILE_ROWS);
    const int64_t tile = gid / TILE_ROWS;
    const int64_t KC = Kpad / CHUNK_K;
    const int64_t rt = tile / KC, kc = tile % KC;
    const int64_t n = rt * TILE_ROWS + r;
    const int64_t nb5 = K / 20 * 4;

    uint32_t words[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int64_t k0 = kc * CHUNK_K + 16 * j;
        uint32_t w = 0;
        if (n < N) {
            if (!I1 && k0 + 16 <= K) {
                // fast path: source byte i = bit pair i, base-3 digit b = byte b
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    uint32_t v = src[gguf_byte_pos(il, k0 / 4 + i, n, N)];
#pragma unroll
                    for (int b = 0; b < 4; b++) { w |= (v % 3u) << (8 * b + 2 * i); v /= 3u; }
                }
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int b = 0; b < 4; b++) {
                        const int64_t k = k0 + 4 * i + b;
                        const uint32_t c = (k < K) ? gguf_code<I1>(src, il, n, k, N, nb5) : 1u;
                        w |= c << (8 * b + 2 * i);
                    }
            }
        } else {
            w = 0x55555555u;  // all codes 1 = weight 0
        }
        words[j] = w;
    }
    reinterpret_cast<uint4 *>(dst)[gid] = make_uint4(words[0], words[1], words[2], words[3]);
}

cudaError_t launch_repack(int type, int64_t K, int64_t N, const uint8_t *gguf, uint8_t *t2, cudaStream_t s) {
    const int64_t Kpad = round_up(K, BLOCK_K), Npad = round_up(N, TILE_ROWS);
    const int64_t total = Npad * Kpad / 4 / 16;
    const int threads = 256;
    const unsigned blocks = (unsigned) ceil_div(total, threads);
    const int il = gguf_interleave(type);
    if (type_is_i1(type))
        repack_kernel<true><<<blocks, threads, 0, s>>>(gguf, t2, il, K, N, Kpad, Npad);
    else
        repack_kernel<false><<<blocks, threads, 0, s>>>(gguf, t2, il, K, N, Kpad, Npad);
    return cudaGetLastError();
}

// T2 -> ternary int8 [N][K] (test hook: round trip of the repacker)
__global__ void unpack_t2_kernel(const uint8_t *__restrict__ t2, int8_t *__restrict__ w, int64_t K, int64_t N,
                                 int64_t Kpad) {
    const int64_t idx = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= N * K) return;
    const int64_t n = idx / K, k = idx % K;
    const uint32_t word = *reinterpret_cast<const uint32_t *>(t2 + t2_word_offset(n, k, Kpad));
    w[idx] = (int8_t) ((int) ((word >> t2_shift(k)) & 3u) - 1);
}

cudaError_t launch_unpack_t2(int64_t K, int64_t N, const uint8_t *t2, int8_t *w, cudaStream_t s) {
    const int64_t Kpad = round_up(K, BLOCK_K);
    const int64_t total = N * K;
    unpack_t2_kernel<<<(unsigned) ceil_div(total, 256), 256, 0, s>>>(t2, w, K, N, Kpad);
    return cudaGetLastError();
}

}  // namespace vlut
