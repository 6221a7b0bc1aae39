/*
 * vlut_b200.h -- C ABI of the B200-native Vec-LUT ternary MUL_MAT path (libvlut_b200.so).
 *
 * This is the drop-in boundary below the ggml backend plugin (libggml-vlut-b200.so):
 * plain pointers and sizes, no C++/torch types, int error codes, never throws.  Every entry point names the
 * piece of the reference (spinal-cord/vlut.cpp, paths relative to its root) it replaces.
 *
 * The op:  out[t][n] = (float)( sum_g wrap16( sum_{k in g} w[n][k] * q[t][k] ) ) * scale[t]
 *   w  ternary {-1,0,+1}, GGUF-packed as I2_V / I2_V_4 / I2_V_8 (4 trits/byte) or I1_V / I1_V_2 (5 trits/byte)
 *   q  per-token int8 quantisation of the f32 activations, scale[t] = max_k(x[t][k]) / 127
 *   g  K-groups of 512 (I2*) or 640 (I1*) inside which the reference keeps int16 partial sums
 * bit-exact against the reference CPU kernels (ggml/src/ggml-cpu/ggml-cpu-quants-vlut.c).
 *
 * There is no CPU fallback: every compute entry point returns VLUT_ERR_CUDA if no sm_100 device is usable.
 */
#ifndef VLUT_B200_H
#define VLUT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VLUT_B200_VERSION 1

/* error codes */
#define VLUT_OK               0
#define VLUT_ERR_INVALID     (-1)  /* bad argument (type, shape, null pointer) */
#define VLUT_ERR_CUDA        (-2)  /* CUDA runtime error or no usable device; see vlut_last_cuda_error() */
#define VLUT_ERR_UNSUPPORTED (-3)  /* shape not handled by the requested algorithm */
#define VLUT_ERR_NOMEM       (-4)

/* weight types: numerically identical to the reference's enum ggml_type (ggml/include/ggml.h:425-434) */
#define VLUT_TYPE_I2_V    40
#define VLUT_TYPE_I2_V_4  42
#define VLUT_TYPE_I2_V_8  43
#define VLUT_TYPE_I1_V    45
#define VLUT_TYPE_I1_V_2  46

/* GEMM algorithms (stage 2).  AUTO picks per token count from the measured crossover table. */
#define VLUT_ALGO_AUTO    0
#define VLUT_ALGO_SIMPLE  1   /* one thread per output, scalar: the in-library correctness anchor            */
#define VLUT_ALGO_LUT     4   /* faithful shared-memory vector-LUT lookup-accumulate, T <= 16 (measured candidate)                  */
#define VLUT_ALGO_TCP     7   /* tcgen05 kind::i8 on CTA pairs (cluster of 2, cta_group::2, M = 256), T > 32: the tensor-bound regime */
#define VLUT_ALGO_TCS     8   /* tcgen05, T <= 64: per-thread cp.async weight stream, K split over a thread-block cluster (DSMEM)    */
#define VLUT_ALGO_GEMV    9   /* CUDA cores, T <= 8: register decode + dp4a, split-K, activation quantisation fused (one launch)     */

typedef struct vlut_ctx vlut_ctx;

/* ---- library / device ---------------------------------------------------------------------------------- */
int         vlut_version(void);
const char *vlut_error_string(int code);
const char *vlut_last_cuda_error(void);            /* text of the last CUDA error seen by this thread       */
int         vlut_device_count(int *count);         /* number of CUDA devices with compute capability 10.x    */
int         vlut_device_info(int device, char *name, size_t name_len, size_t *free_bytes, size_t *total_bytes,
                             int *sm_count);       /* replaces ggml_backend_dev get_description/get_memory    */

/* A context = one device + one stream + a growable device workspace (quantised activations, stream-K
 * partials).  Replaces the per-op scratch of the reference (params->wdata, ggml-cpu.c:1388-1399, and the
 * malloc/free inside every kernel call, ggml-cpu-quants-vlut.c:1177-1182).  `stream` may be NULL (the
 * context creates its own non-blocking stream) or an existing cudaStream_t; name the legacy default stream by
 * CUDA's alias cudaStreamLegacy ((cudaStream_t)0x1).  One host thread per context at a time. */
int   vlut_ctx_create(int device, void *stream, vlut_ctx **out);
void  vlut_ctx_destroy(vlut_ctx *ctx);
void *vlut_ctx_stream(vlut_ctx *ctx);
int   vlut_ctx_sync(vlut_ctx *ctx);
/* Context options.  VLUT_OPT_INPUTS_VISIBLE = 1: the caller promises that the activations it passes are already
 * visible on the device when the call is made -- they were written by a copy, by a kernel that has been synchronised
 * with, or by this context's own kernels -- never by a foreign kernel launched immediately before on the context's
 * stream.  The quantise kernel may then read them ahead of its programmatic-launch dependency wait (i.e. while the
 * previous GEMM of this context is still running).  True for the ggml plugin (src1 arrives through a synchronous
 * set_tensor or from the plugin's own kernels).  Default 0. */
#define VLUT_OPT_INPUTS_VISIBLE 1
int   vlut_ctx_set_option(vlut_ctx *ctx, int option, int value);
/* Re-read the VLUT_* environment knobs (they are read once, at the first call / every vlut_ctx_create -- never on the
 * launch path).  Test hook. */
int   vlut_reload_env(void);
/* number of this library's kernels launched through ctx so far (bench.py's gpu_launches) */
int64_t vlut_ctx_launch_count(const vlut_ctx *ctx);

/* ---- device memory (so that hosts of this ABI -- the ggml backend plugin, FFI bindings -- need no CUDA headers) --
 * Back ggml_backend_buffer_type_i.alloc_buffer / ggml_backend_buffer_i.{free_buffer,set_tensor,get_tensor,
 * memset_tensor,clear} (ggml/src/ggml-backend-impl.h:17-58).  Copies are synchronous on return. */
int vlut_dev_malloc(int device, size_t nbytes, void **out_dev);
int vlut_dev_free(int device, void *dev_ptr);
int vlut_dev_memcpy_h2d(vlut_ctx *ctx, void *dst_dev, const void *src_host, size_t nbytes);
int vlut_dev_memcpy_d2h(vlut_ctx *ctx, void *dst_host, const void *src_dev, size_t nbytes);
int vlut_dev_memset(vlut_ctx *ctx, void *dst_dev, int value, size_t nbytes);

/* ---- weights ------------------------------------------------------------------------------------------- */
/* bytes of the GGUF-packed tensor = ggml_nbytes (ggml/src/ggml.c:1286-1289, :1326-1335); 0 if invalid */
size_t vlut_gguf_nbytes(int type, int64_t K, int64_t N);
/* bytes of the device-resident tiled 2-bit layout ("T2") the kernels stream; what the ggml
 * buffer type's get_alloc_size reports */
size_t vlut_repacked_nbytes(int type, int64_t K, int64_t N);
/* GGUF layout (already on the device) -> T2.  Replaces nothing in the reference (its CPU kernels read the
 * GGUF layout directly); this is the backend buffer's set_tensor hook. */
int vlut_repack_weights(vlut_ctx *ctx, int type, int64_t K, int64_t N, const void *gguf_dev, void *t2_dev);
/* convenience: host GGUF bytes -> staged H2D -> T2 */
int vlut_upload_weights(vlut_ctx *ctx, int type, int64_t K, int64_t N, const void *gguf_host, void *t2_dev);
/* host-side row shard of a GGUF-packed matrix: rows [row0, row0+rows) of all K, in the same packing, for
 * output-row tensor parallelism (the per-thread row split of ggml-cpu.c:1413-1418 generalised to GPUs) */
int vlut_shard_rows_host(int type, int64_t K, int64_t N, int64_t row0, int64_t rows, const void *gguf_host,
                         void *shard_host);

/* ---- the three stages ---------------------------------------------------------------------------------- */
/* (1) per-token int8 activation quantisation into the context workspace.
 *     Replaces quantize_row_i8_v_tile (ggml-cpu-quants-vlut.c:952-973) + the loop at ggml-cpu.c:1394-1407.
 *     x_dev: f32 [T][ldx], ldx >= K. */
int vlut_quantize_act(vlut_ctx *ctx, int type, const float *x_dev, int64_t ldx, int64_t K, int64_t T);
/* (2)+(3) Vec-LUT GEMM + fp32 scale epilogue on the activations last quantised into ctx.
 *     Replaces ggml_gemm_{i2v,i2v4,i2v8,i1v,i1v2}_i8v_lut (ggml-cpu-quants-vlut.c:980-1739).
 *     out_dev: f32 [T][ldo], ldo >= N.  w_scale: optional per-tensor weight scale folded into the epilogue
 *     (the trailing GGML_OP_MUL by the {1} `.scale` tensor, src/models/bitnet.cpp:33-34); pass 1.0f for the
 *     bare op.  With w_scale == 1.0f the result is bit-identical to the reference kernel. */
int vlut_gemm(vlut_ctx *ctx, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev, float *out_dev,
              int64_t ldo, float w_scale, int algo);
/* the whole op on device buffers = GGML_OP_MUL_MAT(vlut weight, f32) of ggml-cpu.c:1372-1421 */
int vlut_mul_mat(vlut_ctx *ctx, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev, const float *x_dev,
                 int64_t ldx, float *out_dev, int64_t ldo, float w_scale, int algo);
/* The whole op for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Needs no vlut_ctx and calls no cudaMalloc / cudaFree / allocating library routine: the quantised activations -- the scratch
 * every other entry point keeps in the context workspace -- live in one file-scope __device__ array of
 * VLUT_NO_ALLOC_MAX_SCRATCH_BYTES that every GPU loading the library holds until the process exits (no_alloc.cu).
 *   limit:   round_up(4 T, 256) + T * round_up(K, 128) <= VLUT_NO_ALLOC_MAX_SCRATCH_BYTES (e.g. 1023 tokens of K = 4096, 292 of
 *            K = 14336) and T <= 65535; beyond it the call returns VLUT_ERR_UNSUPPORTED and says so in `err`.
 *   sharing: every call uses the same array (which keeps the last call's contents; the kernels rewrite every byte they read).
 *            Calls that overlap on the GPU -- different streams -- RACE; calls on one stream are ordered by it.
 * device: CUDA ordinal the buffers live on; stream: cudaStream_t (NULL = default stream); t2_dev: weights repacked by
 * vlut_repack_weights / vlut_upload_weights; x_dev f32 [T][ldx]; out_dev f32 [T][ldo].  err (optional): receives a
 * NUL-terminated message when the result is not VLUT_OK.  Results are bit-identical to vlut_mul_mat's; the kernels are the
 * plain CUDA-core anchor (VLUT_ALGO_SIMPLE class), not the tcgen05 path. */
#define VLUT_NO_ALLOC_MAX_SCRATCH_BYTES ((size_t) 4 << 20)
int vlut_mul_mat_no_alloc(int device, void *stream, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev,
                          const float *x_dev, int64_t ldx, float *out_dev, int64_t ldo, float w_scale, char *err,
                          size_t err_len);
/* out[t][n] = x[t][n] * scale on device buffers (in place allowed): GGML_OP_MUL by a one-element f32 tensor -- the
 * `.scale` of a ternary weight (src/models/bitnet.cpp:33-34, src/llama-graph.cpp:780-783; CPU:
 * ggml_compute_forward_mul, ggml/src/ggml-cpu/binary-ops.cpp:152) -- for the cases where the plugin cannot fold it
 * into the preceding GEMM's epilogue through w_scale (the product has other consumers, or is not adjacent). */
int vlut_scale_rows(vlut_ctx *ctx, const float *x_dev, int64_t ldx, float *out_dev, int64_t ldo, int64_t N, int64_t T,
                    float scale);
/* Multi-GPU, row-sharded weights (output-row tensor parallelism: the thread split of ggml-cpu.c:1413-1418 across
 * GPUs): this rank's [K x N] shard produces columns [col0, col0+N) of the [T][N_total] result.  The GEMM epilogue
 * itself performs the all-gather: it stores every element into each rank's result buffer, either through n_out peer
 * pointers (CUDA P2P / symmetric memory over NVLink 5) or, with multicast != 0 and n_out == 1, through ONE NVSwitch
 * multicast address (multimem.st).  outs_dev[i] must already point at column col0 of rank i's buffer; ldo = N_total.
 * The caller follows with a cross-rank barrier (e.g. a symmetric-memory barrier) before reading. */
int vlut_mul_mat_scatter(vlut_ctx *ctx, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev, const float *x_dev,
                         int64_t ldx, float *const *outs_dev, int n_out, int multicast, int64_t ldo, float w_scale,
                         int algo);
/* The same op with the all-gather DECOUPLED from the GEMM's accumulators: the GEMM stores this rank's slice into its OWN
 * buffer (outs_dev[self]) and stamps every finished tile; a small copy kernel on a second stream, resident beside the
 * GEMM on every SM, forwards each stamped tile to the other ranks' buffers while the GEMM goes on.  Same arguments as
 * vlut_mul_mat_scatter without multicast; `self` = index of this rank's buffer in outs_dev.  Falls back to the fused
 * scatter where the CTA-pair kernel is not the one that runs (T <= 64) or the buffers are not 16-byte aligned. */
int vlut_mul_mat_scatter_push(vlut_ctx *ctx, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev,
                              const float *x_dev, int64_t ldx, float *const *outs_dev, int n_out, int self, int64_t ldo,
                              float w_scale);
/* The cross-rank barrier that publishes a fused all-gather.  flags_dev[i] = rank i's flag array as seen from this device:
 * >= 16 uint32, peer-visible (a symmetric-memory signal pad or a vlut_group allocation), zero-initialised.  Rank `rank`
 * advances its epoch (kept in word [8] of its own array, so the call is CUDA-graph capturable), stores it into slot [rank]
 * of every rank's array (st.release.sys behind a system-scope fence) and waits until all slots of its own array have
 * reached it.  Every rank must call it the same number of times.  Enqueued on the context's stream after the scatter GEMM;
 * replaces the synchronisation of the baseline's separate collective (ncclAllGather). */
int vlut_barrier(vlut_ctx *ctx, void *const *flags_dev, int n_ranks, int rank);
/* the whole op on HOST buffers: H2D of x, the op, D2H of out, stream sync -- what a ggml_backend_sched split
 * costs (ggml/src/ggml-backend.cpp:1443-1460).  Pinned host memory makes the copies asynchronous. */
int vlut_mul_mat_host(vlut_ctx *ctx, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev,
                      const float *x_host, int64_t ldx, float *out_host, int64_t ldo, float w_scale, int algo);

/* Several weights that share ONE activation tensor -- Q/K/V on an attention input, up/gate on an FFN input: consecutive
 * build_lora_mm calls on one tensor (src/models/bitnet.cpp:28-60, src/llama-graph.cpp:668-671, :780-783).  One
 * quantisation and, in the small-token regime (T <= 64), ONE kernel launch that streams all n weights; otherwise the GEMMs
 * run back to back on the one quantisation.  All weights have the same K and the same K-group family (I2* or I1*).
 * Bit-identical to n separate vlut_mul_mat calls. */
int vlut_mul_mat_batch(vlut_ctx *ctx, int n, const int *types, int64_t K, const int64_t *Ns, int64_t T,
                       const void *const *t2_dev, const float *x_dev, int64_t ldx, float *const *outs_dev,
                       const int64_t *ldos, const float *w_scales /* NULL: all 1.0f */, int algo);
/* same, on the activations last quantised into ctx (stage 2+3 only) */
int vlut_gemm_batch(vlut_ctx *ctx, int n, const int *types, int64_t K, const int64_t *Ns, int64_t T,
                    const void *const *t2_dev, float *const *outs_dev, const int64_t *ldos, const float *w_scales, int algo);

/* 1 if vlut_mul_mat handles this op (valid vlut type and packing constraints, at most 64 K-groups), else 0: what the ggml
 * plugin's supports_op asks (ggml_backend_device_i.supports_op, ggml/src/ggml-backend-impl.h:166) */
int vlut_op_supported(int type, int64_t K, int64_t N, int64_t T);

/* which algorithm AUTO resolves to for this shape (also what the last vlut_gemm on ctx used) */
int vlut_pick_algo(int type, int64_t K, int64_t N, int64_t T);
int vlut_ctx_last_algo(const vlut_ctx *ctx);

/* ---- multi-GPU behind the C boundary: ONE process, G GPUs of one box ----------------------------------------------------
 * Output-row tensor parallelism (the thread split of ggml-cpu.c:1413-1419 across GPUs).  Member 0 of a group is the ROOT:
 * activations live there, every member quantises them itself (peer read over NVLink), computes its row shard, and its GEMM
 * epilogue stores the slice straight into the destination buffer(s) -- the gather is fused into the GEMM; ordering between the
 * members' streams is by CUDA events.  What the ggml plugin uses for VLUT_B200_GPUS=G (one logical ggml device owning G GPUs).
 * `devices` may name the same ordinal more than once (several members on one GPU: emulation for tests). */
typedef struct vlut_group vlut_group;
typedef struct vlut_group_weight vlut_group_weight;
int       vlut_group_create(int n_dev, const int *devices, vlut_group **out);
void      vlut_group_destroy(vlut_group *g);
int       vlut_group_size(const vlut_group *g);
vlut_ctx *vlut_group_ctx(vlut_group *g, int member);
/* rows [row0, row0+rows) of an N-row weight owned by `member`: whole 128-row strips, shards differ by at most one strip */
int       vlut_group_rows(const vlut_group *g, int64_t N, int member, int64_t *row0, int64_t *rows);
/* bytes of member 0's repacked shard (what a caller that owns member 0's storage -- the ggml buffer -- must provide) */
size_t    vlut_group_shard0_nbytes(const vlut_group *g, int type, int64_t K, int64_t N);
/* Row-shard a GGUF-packed weight on the host and upload + repack every member's shard.  shard0_dev: caller-owned storage
 * for member 0's shard on the root (the ggml tensor's data), or NULL to let the group allocate it. */
int       vlut_group_upload_weights(vlut_group *g, int type, int64_t K, int64_t N, const void *gguf_host, void *shard0_dev,
                                    vlut_group_weight **out);
void      vlut_group_free_weights(vlut_group *g, vlut_group_weight *w);
/* The op.  x_root f32 [T][ldx] on the root.  outs[j] = base of an f32 [T][ldo] result buffer (peer-accessible device memory);
 * n_dst = 1 gathers into one buffer (typically on the root), n_dst = G all-gathers into a buffer on every member.  On return
 * the root's stream is ordered after every member's stores. */
int       vlut_group_mul_mat(vlut_group *g, const vlut_group_weight *w, int64_t T, const float *x_root, int64_t ldx,
                             float *const *outs, int n_dst, int64_t ldo, float w_scale);
int       vlut_group_sync(vlut_group *g);

/* ---- measurement ------------------------------------------------------------------------------------- */
/* The int8 tensor-pipe peak of the context's device, measured: one CTA per SM issuing tcgen05.mma kind::i8 (M128 N256 K32,
 * A in TMEM like the GEMM kernels) back to back, timed with CUDA events.  *tops = 2 * MAC / s over the whole device in
 * TOP/s, *clk_per_mma (optional) = SM clocks per MMA.  bench.py's int8 roofline denominator. */
int vlut_measure_int8_peak(vlut_ctx *ctx, double *tops, double *clk_per_mma);

/* ---- test hooks ---------------------------------------------------------------------------------------- */
/* copy the quantised activations of the last vlut_quantize_act back: q_host int8 [T][K], scales_host [T] */
int vlut_debug_get_act(vlut_ctx *ctx, int64_t K, int64_t T, int8_t *q_host, float *scales_host);
/* decode a T2 buffer back to ternary int8 [N][K] on the host (round-trip check of the repacker) */
int vlut_debug_unpack_t2(vlut_ctx *ctx, int type, int64_t K, int64_t N, const void *t2_dev, int8_t *w_host);

/* cycle counters of the last tcgen05 GEMM launched with env VLUT_TC_DEBUG=1: [256 CTAs][4 roles][8] int64
 * (roles: TMA producer, MMA issuer, decode warp 0, epilogue warp 0; slot 7 = role lifetime) */
int vlut_debug_tc_counters(long long *out_host, int n_words);

#ifdef __cplusplus
}
#endif
#endif /* VLUT_B200_H */
