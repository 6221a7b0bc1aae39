// c_api.cu -- the extern "C" surface declared in include/vlut_b200.h (libvlut_b200.so).
// Host-side orchestration only: argument checks, workspace management, stage sequencing, algorithm choice.
// No arithmetic happens on the host and there is no CPU fallback: without a usable sm_100 device every compute
// entry point returns VLUT_ERR_CUDA.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "vlut_common.cuh"

using namespace vlut;

static thread_local char g_cuda_err[256] = "";

static int cuda_fail(cudaError_t e, const char *what) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s", what, cudaGetErrorString(e));
    return VLUT_ERR_CUDA;
}
#define CK(call)                                           \
    do {                                                   \
        cudaError_t e__ = (call);                          \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
    } while (0)

// ---- run-time knobs: the environment is read here and nowhere else -------------------------------------------------
static int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}
static Knobs g_knobs;
static bool g_knobs_loaded = false;
namespace vlut {
void knobs_reload() {
    Knobs k{};
    k.algo = env_int("VLUT_B200_ALGO", 0);
    k.host_chunk = env_int("VLUT_HOST_CHUNK", 128);
    if (k.host_chunk < 0) k.host_chunk = 0;
    k.pdl = env_int("VLUT_PDL", 1);
    k.tc_fast = env_int("VLUT_TC_FAST", 1);
    k.tc_smem_kb = env_int("VLUT_TC_SMEM_KB", 200);
    k.tc_stages = env_int("VLUT_TC_STAGES", 0);
    k.tc_grid = env_int("VLUT_TC_GRID", 0);
    k.tc_whole = env_int("VLUT_TC_WHOLE", -1);
    k.tc_split_last = env_int("VLUT_TC_SPLIT_LAST", 0);
    k.tc_coop = env_int("VLUT_TC_COOP", 1);
    k.tc_coop_timeout = env_int("VLUT_TC_COOP_TIMEOUT", 400000);
    k.tc_abl = env_int("VLUT_TC_ABL", 0);
    k.tc_debug = env_int("VLUT_TC_DEBUG", 0);
    k.tc_debug_skip = env_int("VLUT_TC_DEBUG_SKIP", 0);
    k.tc_nt = env_int("VLUT_TC_NT", 0);
    k.tc_astages = env_int("VLUT_TC_ASTAGES", 0);
    k.tc_lean = env_int("VLUT_TC_LEAN", 1);
    k.tc_flat = env_int("VLUT_TC_FLAT", 1);
    k.tc_bulk_out = env_int("VLUT_TC_BULK_OUT", 0);
    k.tc_vec_out = env_int("VLUT_TC_VEC_OUT", 0);
    k.tcs_nas = env_int("VLUT_TCS_NAS", 0);
    k.tcs_split = env_int("VLUT_TCS_SPLIT", 0);
    k.tcs_rows = env_int("VLUT_TCS_ROWS", 0);
    k.tcs_issuers = env_int("VLUT_TCS_ISSUERS", 2) == 1 ? 1 : 2;
    k.tcs_depth = env_int("VLUT_TCS_DEPTH", 3);
    k.tcs_dbg_cta = env_int("VLUT_TCS_DBG_CTA", 0);
    k.tcs_dbg_mask = env_int("VLUT_TCS_DBG_MASK", 15);
    k.tcs_pre = env_int("VLUT_TCS_PRE", -1);
    // activation quantise kernel: 0 register-resident rows (default), 2 two-pass (also the unaligned / K > 16384 fallback)
    k.quant_kernel = env_int("VLUT_QUANT_KERNEL", env_int("VLUT_QUANT_2PASS", 0) ? 2 : 0);
    k.quant_threads = env_int("VLUT_QUANT_THREADS", 0);   // 0 = by shape; 128 / 256 / 512
    k.quant_pdl = env_int("VLUT_QUANT_PDL", -1);
    k.quant_early = env_int("VLUT_QUANT_EARLY", -1);
    k.quant_tpc = env_int("VLUT_QUANT_TPC", 0);
    k.scatter_vec = env_int("VLUT_SCATTER_VEC", 1);
    k.scatter_bulk = env_int("VLUT_SCATTER_BULK", 1);
    k.gemv_split = env_int("VLUT_GEMV_SPLIT", 0);
    k.gemv_max_t = env_int("VLUT_GEMV_MAX_T", -1);
    g_knobs = k;
    g_knobs_loaded = true;
}
const Knobs &knobs() {
    if (!g_knobs_loaded) knobs_reload();
    return g_knobs;
}

// zeroed debug buffer shared by the kernels' optional instrumentation (VLUT_TC_DEBUG=1); word 0 = events to skip
constexpr int DBG_WORDS = 256 * 4 * 8 + 4 * 24 * 2;
static long long *g_dbg_buf = nullptr;
long long *debug_buffer(cudaStream_t s, int skip) {
    if (!g_dbg_buf && cudaMalloc(&g_dbg_buf, DBG_WORDS * sizeof(long long)) != cudaSuccess) return nullptr;
    cudaMemsetAsync(g_dbg_buf, 0, DBG_WORDS * sizeof(long long), s);
    const long long sk = skip;
    cudaMemcpyAsync(g_dbg_buf, &sk, sizeof(sk), cudaMemcpyHostToDevice, s);
    return g_dbg_buf;
}
}  // namespace vlut

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    // grow-only; returns true if (re)allocated
    cudaError_t reserve(size_t n, bool *grew = nullptr) {
        if (grew) *grew = false;
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n + n / 4;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) return e;
        cap = want;
        if (grew) *grew = true;
        return cudaSuccess;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct vlut_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    DevBuf q_rm, q_sw, scales, gpre, nowrap, tile_unsafe, part, counters, acc, stage_x, stage_out, stage_w, tile_done;
    uint32_t *push_flags = nullptr;   // set around a push-mode GEMM launch (vlut_mul_mat_scatter_push)
    cudaStream_t push_stream = nullptr;
    cudaEvent_t push_fork = nullptr, push_join = nullptr;
    ActView act{};
    int64_t act_K = 0, act_T = 0;
    int act_type = 0;
    bool act_rm_valid = false;  // q_rm holds the current quantisation (written by the quantiser or materialised from q_sw)
    int64_t launches = 0;
    int last_algo = 0;
    uint32_t epoch = 0;  // quantisation call counter (ActView.epoch)
    // vlut_mul_mat_host pipeline: one copy stream per PCIe direction and one event pair per token chunk
    int inputs_visible = 0;  // VLUT_OPT_INPUTS_VISIBLE
    // Early-read bookkeeping (VLUT_OPT_INPUTS_VISIBLE).  Every kernel of this library signals launch_dependents at its
    // first instruction, so programmatic-launch chains NEST: a quantise kernel that reads x ahead of its dependency wait
    // may run while ANY kernel launched since the last fully serialising launch (one without the programmatic attribute,
    // or a stream sync) is still storing.  `inflight` holds the output ranges of the GEMMs launched since that point;
    // `inflight_unknown` says some launch since then wrote a range we do not know (scatter epilogue, scale_rows, ...).
    struct Range { const char *lo, *hi; };
    std::vector<Range> inflight;
    bool inflight_unknown = false;
    bool prev_tracked = false;   // the last launch was a quantise / GEMM kernel of this library (programmatic chain)
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
    std::vector<cudaEvent_t> ev_in, ev_done;
};

extern "C" {

int vlut_version(void) { return VLUT_B200_VERSION; }

const char *vlut_error_string(int code) {
    switch (code) {
        case VLUT_OK: return "ok";
        case VLUT_ERR_INVALID: return "invalid argument";
        case VLUT_ERR_CUDA: return "CUDA error / no usable sm_100 device";
        case VLUT_ERR_UNSUPPORTED: return "shape not supported by the requested algorithm";
        case VLUT_ERR_NOMEM: return "out of memory";
        default: return "unknown error";
    }
}

const char *vlut_last_cuda_error(void) { return g_cuda_err; }

int vlut_device_count(int *count) {
    if (!count) return VLUT_ERR_INVALID;
    *count = 0;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
    int ok = 0;
    for (int i = 0; i < n; i++) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, i) == cudaSuccess && major == 10) ok++;
    }
    *count = ok;
    return VLUT_OK;
}

int vlut_device_info(int device, char *name, size_t name_len, size_t *free_bytes, size_t *total_bytes, int *sm_count) {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (name && name_len) snprintf(name, name_len, "%s", prop.name);
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (free_bytes || total_bytes) {
        int prev = 0;
        cudaGetDevice(&prev);
        CK(cudaSetDevice(device));
        size_t f = 0, t = 0;
        CK(cudaMemGetInfo(&f, &t));
        if (free_bytes) *free_bytes = f;
        if (total_bytes) *total_bytes = t;
        cudaSetDevice(prev);
    }
    return VLUT_OK;
}

int vlut_reload_env(void) {
    knobs_reload();
    return VLUT_OK;
}

int vlut_ctx_create(int device, void *stream, vlut_ctx **out) {
    if (!out) return VLUT_ERR_INVALID;
    *out = nullptr;
    knobs_reload();
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return VLUT_ERR_INVALID;
    int major = 0;
    CK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "device %d has compute capability %d.x; this library is sm_100a only", device,
                 major);
        return VLUT_ERR_CUDA;
    }
    CK(cudaSetDevice(device));
    vlut_ctx *c = new (std::nothrow) vlut_ctx();
    if (!c) return VLUT_ERR_NOMEM;
    c->device = device;
    CK(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
    if (stream) {
        c->stream = (cudaStream_t) stream;
    } else {
        cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) {
            delete c;
            return cuda_fail(e, "cudaStreamCreate");
        }
        c->own_stream = true;
    }
    *out = c;
    return VLUT_OK;
}

void vlut_ctx_destroy(vlut_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    DevBuf *bufs[] = {&c->q_rm, &c->q_sw,     &c->scales,  &c->gpre,      &c->nowrap, &c->tile_unsafe,
                      &c->part, &c->counters, &c->acc, &c->stage_x, &c->stage_out, &c->stage_w};
    for (DevBuf *b : bufs) b->release();
    for (cudaEvent_t e : c->ev_in) cudaEventDestroy(e);
    for (cudaEvent_t e : c->ev_done) cudaEventDestroy(e);
    if (c->h2d_stream) cudaStreamDestroy(c->h2d_stream);
    if (c->d2h_stream) cudaStreamDestroy(c->d2h_stream);
    if (c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}

void *vlut_ctx_stream(vlut_ctx *c) { return c ? (void *) c->stream : nullptr; }

int vlut_ctx_sync(vlut_ctx *c) {
    if (!c) return VLUT_ERR_INVALID;
    CK(cudaStreamSynchronize(c->stream));
    c->inflight.clear();
    c->inflight_unknown = false;
    return VLUT_OK;
}

int64_t vlut_ctx_launch_count(const vlut_ctx *c) { return c ? c->launches : 0; }
int vlut_ctx_last_algo(const vlut_ctx *c) { return c ? c->last_algo : 0; }

int vlut_dev_malloc(int device, size_t nbytes, void **out_dev) {
    if (!out_dev) return VLUT_ERR_INVALID;
    *out_dev = nullptr;
    CK(cudaSetDevice(device));
    cudaError_t e = cudaMalloc(out_dev, nbytes ? nbytes : 1);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        return VLUT_ERR_NOMEM;
    }
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    return VLUT_OK;
}

int vlut_dev_free(int device, void *dev_ptr) {
    if (!dev_ptr) return VLUT_OK;
    CK(cudaSetDevice(device));
    CK(cudaFree(dev_ptr));
    return VLUT_OK;
}

int vlut_dev_memcpy_h2d(vlut_ctx *c, void *dst_dev, const void *src_host, size_t nbytes) {
    if (!c || (nbytes && (!dst_dev || !src_host))) return VLUT_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(dst_dev, src_host, nbytes, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return VLUT_OK;
}

int vlut_dev_memcpy_d2h(vlut_ctx *c, void *dst_host, const void *src_dev, size_t nbytes) {
    if (!c || (nbytes && (!dst_host || !src_dev))) return VLUT_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    CK(cudaMemcpyAsync(dst_host, src_dev, nbytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return VLUT_OK;
}

int vlut_dev_memset(vlut_ctx *c, void *dst_dev, int value, size_t nbytes) {
    if (!c || (nbytes && !dst_dev)) return VLUT_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    CK(cudaMemsetAsync(dst_dev, value, nbytes, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return VLUT_OK;
}

static int check_shape(int type, int64_t K, int64_t N);
// A kernel of ours that is launched WITHOUT the programmatic attribute and never signals launch_dependents (repack,
// scale_rows, the anchor / LUT kernels, re-tiling, the barrier, the peak loop) was just launched.  It starts only after
// everything before it has completed, and -- its completion being the implicit trigger -- its successor starts only after it
// has completed: a full serialisation point of the stream.  Nothing launched before or by it can race with a later early read.
static void note_foreign_write(vlut_ctx *c) {
    c->inflight.clear();
    c->inflight_unknown = false;
    c->prev_tracked = false;
}

size_t vlut_gguf_nbytes(int type, int64_t K, int64_t N) {
    if (check_shape(type, K, N) != VLUT_OK) return 0;
    return (size_t) (gguf_row_bytes(type, K) * N);
}

size_t vlut_repacked_nbytes(int type, int64_t K, int64_t N) {
    if (check_shape(type, K, N) != VLUT_OK) return 0;
    return (size_t) (round_up(N, TILE_ROWS) * round_up(K, BLOCK_K) / 4);
}

static int check_shape(int type, int64_t K, int64_t N) {
    if (!type_valid(type) || K <= 0 || N <= 0 || K % 4) return VLUT_ERR_INVALID;
    // the reference packers index with int and require these (ggml-quants-vlut.c:61-107): K%16 for I2_V_4, K%32 for I2_V_8
    if (type == VLUT_TYPE_I2_V_4 && K % 16) return VLUT_ERR_INVALID;
    if (type == VLUT_TYPE_I2_V_8 && K % 32) return VLUT_ERR_INVALID;
    return VLUT_OK;
}

int vlut_repack_weights(vlut_ctx *c, int type, int64_t K, int64_t N, const void *gguf_dev, void *t2_dev) {
    if (!c || !gguf_dev || !t2_dev) return VLUT_ERR_INVALID;
    int rc = check_shape(type, K, N);
    if (rc) return rc;
    CK(cudaSetDevice(c->device));
    note_foreign_write(c);
    CK(launch_repack(type, K, N, (const uint8_t *) gguf_dev, (uint8_t *) t2_dev, c->stream));
    c->launches++;
    return VLUT_OK;
}

int vlut_upload_weights(vlut_ctx *c, int type, int64_t K, int64_t N, const void *gguf_host, void *t2_dev) {
    if (!c || !gguf_host || !t2_dev) return VLUT_ERR_INVALID;
    int rc = check_shape(type, K, N);
    if (rc) return rc;
    CK(cudaSetDevice(c->device));
    const size_t nb = vlut_gguf_nbytes(type, K, N);
    // the staging buffer may still be read by a previous repack on this stream
    CK(cudaStreamSynchronize(c->stream));
    if (c->stage_w.reserve(nb) != cudaSuccess) return VLUT_ERR_NOMEM;
    CK(cudaMemcpyAsync(c->stage_w.p, gguf_host, nb, cudaMemcpyHostToDevice, c->stream));
    CK(launch_repack(type, K, N, (const uint8_t *) c->stage_w.p, (uint8_t *) t2_dev, c->stream));
    c->launches++;
    CK(cudaStreamSynchronize(c->stream));
    return VLUT_OK;
}

int vlut_shard_rows_host(int type, int64_t K, int64_t N, int64_t row0, int64_t rows, const void *gguf_host,
                         void *shard_host) {
    if (!gguf_host || !shard_host || row0 < 0 || rows <= 0 || row0 + rows > N) return VLUT_ERR_INVALID;
    int rc = check_shape(type, K, N);
    if (rc) return rc;
    const uint8_t *src = (const uint8_t *) gguf_host;
    uint8_t *dst = (uint8_t *) shard_host;
    const int64_t rs = gguf_row_bytes(type, K);
    const int64_t nb5 = type_is_i1(type) ? K / 20 * 4 : rs;
    const int il0 = gguf_interleave(type);
    // byte p of row n lives at (p/il)*il*N + n*il + p%il: for a fixed interleave group the rows are contiguous
    for (int64_t p0 = 0; p0 < rs;) {
        const int il = (type == VLUT_TYPE_I1_V_2 && p0 >= nb5) ? 1 : il0;
        memcpy(dst + (p0 / il) * il * rows, src + (p0 / il) * il * N + row0 * il, (size_t) (rows * il));
        p0 += il;
    }
    return VLUT_OK;
}

// (re)shape the activation workspace of ctx for (type, K, T)
// q_sw layout a GEMM kernel wants (ActView.sw_tile / ActView.perm)
static int want_sw_tile(int algo, int NT) { return (algo == VLUT_ALGO_TCP && NT >= 64) ? NT / 2 : NT; }
static int want_perm(int algo, int NT) { return (algo == VLUT_ALGO_TCS && NT <= 32) ? 1 : 0; }

static int prepare_act(vlut_ctx *c, int type, int64_t K, int64_t T, int algo, int64_t N_hint) {
    const int64_t Kpad = round_up(K, BLOCK_K);
    const int NT = (algo == VLUT_ALGO_TCP) ? token_tile_pair(K, N_hint, T, group_size(type), c->sm_count, knobs().tc_nt)
                                           : token_tile(T);
    const int64_t Tpad = round_up(T, NT);
    const int gsize = group_size(type);
    const int G = (int) ceil_div(K, gsize);
    if (G > 64) return VLUT_ERR_UNSUPPORTED;
    if (c->q_rm.reserve((size_t) (Tpad * Kpad)) != cudaSuccess) return VLUT_ERR_NOMEM;
    if (c->q_sw.reserve((size_t) (Tpad * Kpad)) != cudaSuccess) return VLUT_ERR_NOMEM;
    if (c->scales.reserve((size_t) Tpad * sizeof(float)) != cudaSuccess) return VLUT_ERR_NOMEM;
    if (c->gpre.reserve((size_t) Tpad * (G + 1) * sizeof(int32_t)) != cudaSuccess) return VLUT_ERR_NOMEM;
    if (c->nowrap.reserve((size_t) Tpad * sizeof(int32_t)) != cudaSuccess) return VLUT_ERR_NOMEM;
    bool grew = false;
    if (c->tile_unsafe.reserve((size_t) (Tpad / 16 + 1) * sizeof(uint32_t), &grew) != cudaSuccess) return VLUT_ERR_NOMEM;
    if (grew && cudaMemsetAsync(c->tile_unsafe.p, 0, c->tile_unsafe.cap, c->stream) != cudaSuccess) return VLUT_ERR_CUDA;
    c->act.q_rm = (int8_t *) c->q_rm.p;
    c->act.q_sw = (int8_t *) c->q_sw.p;
    c->act.scales = (float *) c->scales.p;
    c->act.gpre = (int32_t *) c->gpre.p;
    c->act.nowrap = (int32_t *) c->nowrap.p;
    c->act.tile_unsafe = (uint32_t *) c->tile_unsafe.p;
    if (++c->epoch == 0) c->epoch = 1;
    c->act.epoch = c->epoch;
    c->act.Kpad = Kpad;
    c->act.Tpad = Tpad;
    c->act.G = G;
    c->act.gsize = gsize;
    c->act.NT = NT;
    c->act.sw_tile = want_sw_tile(algo, NT);
    c->act.perm = want_perm(algo, NT);
    c->act_K = K;
    c->act_T = T;
    c->act_type = type;
    return VLUT_OK;
}

// Launch policy of the quantise kernel (measured on the quantise + GEMM chain at 4096 -> 14336, scripts/time_op.py):
//   pdl    programmatic launch behind the previous GEMM: gains 0.1-0.5 us for T <= 64 and from T = 1024 up, loses 1.6-2 us
//          in between
//   early  programmatic launch with the row loads, maximum and scale AHEAD of the dependency wait: only where the host can
//          prove that no kernel still in flight writes x (early_ok) and where it measured faster (T <= 64, 512 <= T <= 1024)
struct QuantPolicy { bool pdl, early; };
static QuantPolicy quant_policy(int64_t T, bool early_ok, bool opted_in) {
    const Knobs &k = knobs();
    QuantPolicy q{false, false};
    if (!k.pdl) return q;
    if (early_ok) {
        if (k.quant_early >= 0) q.early = k.quant_early != 0;    // VLUT_QUANT_EARLY=1: experiment without the option
        else q.early = opted_in && (T <= 64 || (T >= 512 && T <= 1024));
    }
    q.pdl = q.early || (k.quant_pdl >= 0 ? k.quant_pdl != 0 : (T <= 64 || T > 512));
    return q;
}

// quantise for the kernel `algo` resolves to (the CTA-pair kernel wants half-height swizzled tiles)
static int quantize_impl(vlut_ctx *c, int type, const float *x_dev, int64_t ldx, int64_t K, int64_t T, int algo, int64_t N_hint) {
    if (!c || !x_dev || T <= 0 || ldx < K) return VLUT_ERR_INVALID;
    if (!type_valid(type) || K <= 0 || K % 4) return VLUT_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    if (algo == VLUT_ALGO_AUTO) algo = vlut_pick_algo(type, K, N_hint, T);
    int rc = prepare_act(c, type, K, T, algo, N_hint);
    if (rc) return rc;
    // x is not written by any kernel of this context that may still be running: the last launch was a GEMM and the
    // output ranges of every GEMM launched since the last fully serialising launch are known and do not meet x
    const char *xl = (const char *) x_dev, *xh = xl + ((size_t) (T - 1) * ldx + K) * sizeof(float);
    // ... and the launch right before this one is one of the tracked kernels (quantise / GEMM): behind a plain launch the
    // early reads would rely on the implicit completion trigger alone, without the dependency wait's memory flush
    bool untouched = c->prev_tracked && !c->inflight_unknown;
    for (const vlut_ctx::Range &r : c->inflight) untouched = untouched && (xh <= r.lo || xl >= r.hi);
    const QuantPolicy qp = quant_policy(T, untouched, c->inputs_visible != 0);
    // q_rm (row-major copy) is only written for the kernels that read it; otherwise it is materialised on demand
    const bool want_rm = (algo == VLUT_ALGO_SIMPLE || algo == VLUT_ALGO_LUT);
    CK(launch_quantize(x_dev, ldx, K, T, c->act, c->stream, qp.pdl, qp.early, want_rm));
    c->act_rm_valid = want_rm;
    c->launches++;
    c->prev_tracked = true;
    if (!qp.pdl) {   // launched without the programmatic attribute: starts after everything before it has completed
        c->inflight.clear();
        c->inflight_unknown = false;
    }
    return VLUT_OK;
}

// the row-major copy of the quantised activations (anchor kernel, LUT candidate, re-tiling, test read-back)
static int ensure_q_rm(vlut_ctx *c) {
    if (c->act_rm_valid) return VLUT_OK;
    CK(launch_unswizzle(c->act, c->stream));
    c->launches++;
    c->act_rm_valid = true;
    c->inflight.clear();           // plain launch: serialises the stream
    c->inflight_unknown = false;
    c->prev_tracked = false;
    return VLUT_OK;
}

int vlut_quantize_act(vlut_ctx *c, int type, const float *x_dev, int64_t ldx, int64_t K, int64_t T) {
    // the weight's row count is unknown here: the cost model then picks the 128-token tile for the CTA-pair kernel (the
    // faster one for most shapes; vlut_mul_mat / vlut_mul_mat_batch quantise with the real row count)
    return quantize_impl(c, type, x_dev, ldx, K, T, VLUT_ALGO_AUTO, 1);
}

int vlut_op_supported(int type, int64_t K, int64_t N, int64_t T) {
    if (check_shape(type, K, N) != VLUT_OK || T <= 0) return 0;
    if (ceil_div(K, group_size(type)) > 64) return 0;   // prepare_act: at most 64 K-groups (K <= 32768 I2*, 40960 I1*)
    return 1;
}

// AUTO runs the CUDA-core GEMV (quantisation fused, one launch) up to this many tokens (measured crossover)
constexpr int GEMV_MAX_T_DEFAULT = 0;

int vlut_pick_algo(int type, int64_t K, int64_t N, int64_t T) {
    const int forced = knobs().algo;
    if (forced > 0) return forced;
    // crossover table from B200 measurements (profiles/):
    //   T <= 4    GEMV (CUDA cores: register decode + dp4a, split-K, quantisation fused) where it measured faster
    //   T <= 64   TCS  (per-thread cp.async weight stream, 16 decode warps, two MMA issuers, cluster split-K)
    //   T  > 64   TCP  (CTA pairs, cta_group::2: each CTA stages half of the token tile -> half the activation fill)
    const int gemv_max = knobs().gemv_max_t >= 0 ? knobs().gemv_max_t : GEMV_MAX_T_DEFAULT;
    if (T <= gemv_max && T <= 8 && vlut_op_supported(type, K, N, T)) return VLUT_ALGO_GEMV;
    if (T <= 64 && gemm_tcs_supported(type, K, N, T)) return VLUT_ALGO_TCS;
    if (T > 64 && gemm_tc5_supported(type, K, N, T)) return VLUT_ALGO_TCP;
    return VLUT_ALGO_SIMPLE;
}

// One launch over up to MAX_BATCH weights that share the quantised activations (Q/K/V on one attention input, up/gate on
// one FFN input: consecutive build_lora_mm calls on one tensor, src/models/bitnet.cpp:28-60, src/llama-graph.cpp:668-671).
struct GemmCall {
    int type;
    int64_t N;
    const void *t2;
    float *out;
    int64_t ldo;
    float w_scale;
};

static int gemm_impl(vlut_ctx *c, int64_t K, int64_t T, const GemmCall *calls, int ncalls, float *const *outs, int n_out,
                     int multicast, int algo) {
    if (!c || !calls || ncalls < 1 || ncalls > MAX_BATCH || T <= 0 || n_out > 8) return VLUT_ERR_INVALID;
    if (ncalls > 1 && n_out > 0) return VLUT_ERR_UNSUPPORTED;
    int64_t Nsum = 0;
    for (int i = 0; i < ncalls; i++) {
        const GemmCall &q = calls[i];
        if (!q.t2 || (!q.out && n_out <= 0) || q.ldo < q.N) return VLUT_ERR_INVALID;
        int rc = check_shape(q.type, K, q.N);
        if (rc) return rc;
        if (group_size(c->act_type) != group_size(q.type)) return VLUT_ERR_INVALID;
        Nsum += round_up(q.N, TILE_ROWS);
    }
    if (c->act_K != K || c->act_T != T) return VLUT_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    if (algo == VLUT_ALGO_AUTO) {
        algo = vlut_pick_algo(calls[0].type, K, Nsum, T);
        if (algo == VLUT_ALGO_GEMV) algo = VLUT_ALGO_TCS;   // stage 2 alone: the GEMV kernel is the whole op (it quantises x itself)
    }
    if (algo == VLUT_ALGO_GEMV) return VLUT_ERR_UNSUPPORTED;
    // only the small-T streaming kernel walks several weights in one launch; everything else runs them back to back
    if (ncalls > 1 && algo != VLUT_ALGO_TCS) {
        for (int i = 0; i < ncalls; i++) {
            int rc = gemm_impl(c, K, T, calls + i, 1, nullptr, 0, 0, algo);
            if (rc) return rc;
        }
        return VLUT_OK;
    }

    GemmArgs g{};
    g.act = c->act;
    g.n_out = n_out > 0 ? n_out : 0;
    g.multicast = multicast;
    g.vec_ok = (g.n_out > 0 && calls[0].ldo % 4 == 0) ? 1 : 0;
    for (int i = 0; i < g.n_out; i++) {
        g.outs[i] = outs[i];
        if (reinterpret_cast<uintptr_t>(outs[i]) & 15) g.vec_ok = 0;
    }
    if (knobs().scatter_vec == 0) g.vec_ok = 0;
    g.bulk_ok = (g.vec_ok && !multicast && calls[0].N % 4 == 0 && knobs().scatter_bulk) ? 1 : 0;
    g.K = K;
    g.T = T;
    g.sm_count = c->sm_count;
    g.device = c->device;
    g.n_tens = ncalls;
    int strips = 0;
    for (int i = 0; i < ncalls; i++) {
        TensorRef &t = g.tens[i];
        t.w = (const uint8_t *) calls[i].t2;
        t.out = calls[i].out;
        t.N = calls[i].N;
        t.ldo = calls[i].ldo;
        t.w_scale = calls[i].w_scale;
        t.strip0 = strips;
        strips += (int) (round_up(t.N, TILE_ROWS) / TILE_ROWS);
    }
    g.tens[ncalls].strip0 = strips;   // sentinel
    // single-tensor view (every kernel but TCS reads these)
    g.w = g.tens[0].w;
    g.out = g.tens[0].out;
    g.ldo = g.tens[0].ldo;
    g.tile_done = c->push_flags;
    g.tile_epoch = c->act.epoch;
    g.local_bulk = (g.n_out == 0 && knobs().tc_bulk_out && g.ldo % 4 == 0 && calls[0].N % 4 == 0 &&
                    !(reinterpret_cast<uintptr_t>(g.out) & 15)) ? 1 : 0;
    g.N = g.tens[0].N;
    g.Npad = round_up(g.N, TILE_ROWS);
    g.type = calls[0].type;
    g.w_scale = g.tens[0].w_scale;

    if (algo == VLUT_ALGO_TCP || algo == VLUT_ALGO_LUT) {
        // stream-K workspace: private partial-tile slots per persistent CTA (no initialisation needed), self-resetting
        // per-tile arrival counters and accumulation tiles (zeroed once at (re)allocation)
        bool grew = false;
        const size_t pb = (size_t) c->sm_count * 6 * 128 * 128 * sizeof(int32_t);
        if (c->part.reserve(pb) != cudaSuccess) return VLUT_ERR_NOMEM;
        const size_t cb = (size_t) ((g.Npad / TILE_ROWS + 1) * (c->act.Tpad / c->act.NT)) * sizeof(int32_t);
        if (c->counters.reserve(cb, &grew) != cudaSuccess) return VLUT_ERR_NOMEM;
        if (grew) CK(cudaMemsetAsync(c->counters.p, 0, c->counters.cap, c->stream));
        const size_t ab = (size_t) c->sm_count * 128 * 256 * sizeof(int32_t);
        if (c->acc.reserve(ab, &grew) != cudaSuccess) return VLUT_ERR_NOMEM;
        if (grew) CK(cudaMemsetAsync(c->acc.p, 0, c->acc.cap, c->stream));
        g.acc = (int32_t *) c->acc.p;
        g.part = (int32_t *) c->part.p;
        g.counters = (int32_t *) c->counters.p;
    }

    int launches = 0;
    bool plain_launch = false;   // some kernel went out without the programmatic attribute (serialises the stream)
    cudaError_t e;
    if (algo == VLUT_ALGO_TCP || algo == VLUT_ALGO_TCS) {
        // the workspace was quantised for another kernel's swizzled-tile layout: rewrite q_sw (through q_rm)
        const int want = want_sw_tile(algo, c->act.NT), wantp = want_perm(algo, c->act.NT);
        if (c->act.sw_tile != want || c->act.perm != wantp) {
            int rc = ensure_q_rm(c);
            if (rc) return rc;
            c->act.sw_tile = want;
            c->act.perm = wantp;
            CK(launch_relayout(c->act, c->stream));
            launches++;
            plain_launch = true;
        }
        g.act = c->act;
    } else if (algo == VLUT_ALGO_SIMPLE || algo == VLUT_ALGO_LUT) {
        int rc = ensure_q_rm(c);
        if (rc) return rc;
        plain_launch = true;
    }
    switch (algo) {
        case VLUT_ALGO_SIMPLE:
            if (g.n_out) return VLUT_ERR_UNSUPPORTED;
            e = launch_gemm_simple(g, c->stream);
            launches += 1;
            break;
        case VLUT_ALGO_TCS:
            if (!gemm_tcs_supported(g.type, K, g.N, T)) return VLUT_ERR_UNSUPPORTED;
            e = launch_gemm_tcs(g, c->stream, &launches);
            break;
        case VLUT_ALGO_TCP:
            if (!gemm_tc5_supported(g.type, K, g.N, T)) return VLUT_ERR_UNSUPPORTED;
            e = launch_gemm_tc5(g, c->stream, &launches);
            break;
        case VLUT_ALGO_LUT:
            if (!gemm_lut_supported(g.type, K, g.N, T) || g.n_out) return VLUT_ERR_UNSUPPORTED;
            // the LUT candidate accumulates with atomics into the (shared) workspace: hand it a zeroed region
            if ((size_t) (16 * g.Npad) * sizeof(int32_t) > c->part.cap) return VLUT_ERR_UNSUPPORTED;
            CK(cudaMemsetAsync(c->part.p, 0, (size_t) (16 * g.Npad) * sizeof(int32_t), c->stream));
            e = launch_gemm_lut(g, c->stream, &launches);
            break;
        default:
            return VLUT_ERR_INVALID;
    }
    if (e != cudaSuccess) return cuda_fail(e, "gemm launch");
    c->launches += launches;
    c->last_algo = algo;
    // early-read bookkeeping: the output ranges now in flight
    if (plain_launch || !knobs().pdl || algo == VLUT_ALGO_SIMPLE || algo == VLUT_ALGO_LUT) {
        c->inflight.clear();
        c->inflight_unknown = false;
    }
    c->prev_tracked = knobs().pdl && algo != VLUT_ALGO_SIMPLE && algo != VLUT_ALGO_LUT;
    if (n_out > 0 || c->inflight.size() >= 8) {
        c->inflight_unknown = true;
    } else {
        for (int i = 0; i < ncalls; i++) {
            const char *lo = (const char *) calls[i].out;
            c->inflight.push_back({lo, lo + ((size_t) (T - 1) * calls[i].ldo + calls[i].N) * sizeof(float)});
        }
    }
    return VLUT_OK;
}

// the whole op in ONE launch on the CUDA cores (quantisation fused): does not touch the activation workspace
static int gemv_impl(vlut_ctx *c, int64_t K, int64_t T, const GemmCall *calls, int ncalls, const float *x_dev, int64_t ldx) {
    if (!c || !calls || !x_dev || ncalls < 1 || ncalls > MAX_BATCH || T <= 0 || ldx < K) return VLUT_ERR_INVALID;
    GemmArgs g{};
    g.K = K;
    g.T = T;
    g.sm_count = c->sm_count;
    g.device = c->device;
    g.n_tens = ncalls;
    int strips = 0;
    for (int i = 0; i < ncalls; i++) {
        const GemmCall &q = calls[i];
        if (!q.t2 || !q.out || q.ldo < q.N) return VLUT_ERR_INVALID;
        int rc = check_shape(q.type, K, q.N);
        if (rc) return rc;
        if (group_size(q.type) != group_size(calls[0].type)) return VLUT_ERR_INVALID;
        g.tens[i] = TensorRef{(const uint8_t *) q.t2, q.out, q.N, q.ldo, q.w_scale, strips};
        strips += (int) (round_up(q.N, TILE_ROWS) / TILE_ROWS);
    }
    g.tens[ncalls].strip0 = strips;
    g.type = calls[0].type;
    g.N = calls[0].N;
    if (!gemm_gemv_supported(g.type, K, g.N, T, x_dev, ldx)) return VLUT_ERR_UNSUPPORTED;
    CK(cudaSetDevice(c->device));
    int launches = 0;
    cudaError_t e = launch_gemm_gemv(g, x_dev, ldx, c->stream, &launches);
    if (e != cudaSuccess) return cuda_fail(e, "gemv launch");
    c->launches += launches;
    c->last_algo = VLUT_ALGO_GEMV;
    c->act_T = 0;   // the workspace holds no quantisation of this x
    c->prev_tracked = true;
    if (!knobs().pdl) {
        c->inflight.clear();
        c->inflight_unknown = false;
    }
    if (c->inflight.size() >= 8) {
        c->inflight_unknown = true;
    } else {
        for (int i = 0; i < ncalls; i++) {
            const char *lo = (const char *) calls[i].out;
            c->inflight.push_back({lo, lo + ((size_t) (T - 1) * calls[i].ldo + calls[i].N) * sizeof(float)});
        }
    }
    return VLUT_OK;
}

// AUTO -> GEMV only where the kernel can run (aligned rows); a forced GEMV that cannot run reports UNSUPPORTED
static bool use_gemv(int *algo, int type, int64_t K, int64_t Nsum, int64_t T, const float *x_dev, int64_t ldx) {
    const bool was_auto = (*algo == VLUT_ALGO_AUTO);
    const int a = was_auto ? vlut_pick_algo(type, K, Nsum, T) : *algo;
    if (a != VLUT_ALGO_GEMV) return false;
    if (was_auto && !gemm_gemv_supported(type, K, Nsum, T, x_dev, ldx)) {
        *algo = VLUT_ALGO_TCS;
        return false;
    }
    return true;
}

int vlut_gemm(vlut_ctx *c, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev, float *out_dev, int64_t ldo,
              float w_scale, int algo) {
    if (!out_dev) return VLUT_ERR_INVALID;
    const GemmCall q{type, N, t2_dev, out_dev, ldo, w_scale};
    return gemm_impl(c, K, T, &q, 1, nullptr, 0, 0, algo);
}

int vlut_gemm_batch(vlut_ctx *c, int n, const int *types, int64_t K, const int64_t *Ns, int64_t T, const void *const *t2_dev,
                    float *const *outs_dev, const int64_t *ldos, const float *w_scales, int algo) {
    if (!c || n < 1 || !types || !Ns || !t2_dev || !outs_dev || !ldos) return VLUT_ERR_INVALID;
    for (int i0 = 0; i0 < n; i0 += MAX_BATCH) {
        GemmCall q[MAX_BATCH];
        const int m = std::min(MAX_BATCH, n - i0);
        for (int i = 0; i < m; i++) {
            if (!outs_dev[i0 + i]) return VLUT_ERR_INVALID;
            q[i] = GemmCall{types[i0 + i], Ns[i0 + i], t2_dev[i0 + i], outs_dev[i0 + i], ldos[i0 + i],
                            w_scales ? w_scales[i0 + i] : 1.0f};
        }
        int rc = gemm_impl(c, K, T, q, m, nullptr, 0, 0, algo);
        if (rc) return rc;
    }
    return VLUT_OK;
}

int vlut_mul_mat_batch(vlut_ctx *c, int n, const int *types, int64_t K, const int64_t *Ns, int64_t T, const void *const *t2_dev,
                       const float *x_dev, int64_t ldx, float *const *outs_dev, const int64_t *ldos, const float *w_scales,
                       int algo) {
    if (!c || n < 1 || !types || !Ns) return VLUT_ERR_INVALID;
    if (!t2_dev || !outs_dev || !ldos) return VLUT_ERR_INVALID;
    int64_t Nsum = 0;
    for (int i = 0; i < n; i++) Nsum += Ns[i];
    if (use_gemv(&algo, types[0], K, Nsum, T, x_dev, ldx)) {
        for (int i0 = 0; i0 < n; i0 += MAX_BATCH) {
            GemmCall q[MAX_BATCH];
            const int m = std::min(MAX_BATCH, n - i0);
            for (int i = 0; i < m; i++)
                q[i] = GemmCall{types[i0 + i], Ns[i0 + i], t2_dev[i0 + i], outs_dev[i0 + i], ldos[i0 + i],
                                w_scales ? w_scales[i0 + i] : 1.0f};
            int rc = gemv_impl(c, K, T, q, m, x_dev, ldx);
            if (rc) return rc;
        }
        return VLUT_OK;
    }
    int rc = quantize_impl(c, types[0], x_dev, ldx, K, T, algo, Nsum);
    if (rc) return rc;
    return vlut_gemm_batch(c, n, types, K, Ns, T, t2_dev, outs_dev, ldos, w_scales, algo);
}

int vlut_mul_mat_scatter(vlut_ctx *c, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev, const float *x_dev,
                         int64_t ldx, float *const *outs_dev, int n_out, int multicast, int64_t ldo, float w_scale,
                         int algo) {
    if (!outs_dev || n_out < 1 || n_out > 8 || (multicast && n_out != 1)) return VLUT_ERR_INVALID;
    int rc = quantize_impl(c, type, x_dev, ldx, K, T, algo, N);
    if (rc) return rc;
    const GemmCall q{type, N, t2_dev, nullptr, ldo, w_scale};
    return gemm_impl(c, K, T, &q, 1, outs_dev, n_out, multicast, algo);
}

int vlut_mul_mat_scatter_push(vlut_ctx *c, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev, const float *x_dev,
                              int64_t ldx, float *const *outs_dev, int n_out, int self, int64_t ldo, float w_scale) {
    if (!c || !outs_dev || n_out < 2 || n_out > 8 || self < 0 || self >= n_out) return VLUT_ERR_INVALID;
    // the copy kernel moves 16-byte words of rows that start 16-byte aligned; the CTA-pair kernel stamps the tiles
    bool ok = (ldo % 4 == 0) && (N % 4 == 0) && vlut_pick_algo(type, K, N, T) == VLUT_ALGO_TCP && knobs().algo == 0;
    for (int i = 0; i < n_out; i++) ok = ok && !(reinterpret_cast<uintptr_t>(outs_dev[i]) & 15);
    if (!ok) return vlut_mul_mat_scatter(c, type, K, N, T, t2_dev, x_dev, ldx, outs_dev, n_out, 0, ldo, w_scale, VLUT_ALGO_AUTO);
    int rc = quantize_impl(c, type, x_dev, ldx, K, T, VLUT_ALGO_TCP, N);
    if (rc) return rc;
    CK(cudaSetDevice(c->device));
    if (!c->push_stream) {
        CK(cudaStreamCreateWithFlags(&c->push_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&c->push_fork, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&c->push_join, cudaEventDisableTiming));
    }
    const int NT = c->act.NT, TT = (int) (c->act.Tpad / NT);
    const int RP = (int) ((ceil_div(N, TILE_ROWS) + 1) / 2), ntiles = RP * TT;
    bool grew = false;
    if (c->tile_done.reserve((size_t) ntiles * 2 * sizeof(uint32_t), &grew) != cudaSuccess) return VLUT_ERR_NOMEM;
    if (grew) CK(cudaMemsetAsync(c->tile_done.p, 0, c->tile_done.cap, c->stream));
    // fork: the copy kernel starts with the GEMM (both behind the quantise kernel)
    CK(cudaEventRecord(c->push_fork, c->stream));
    CK(cudaStreamWaitEvent(c->push_stream, c->push_fork, 0));
    c->push_flags = (uint32_t *) c->tile_done.p;
    const GemmCall q{type, N, t2_dev, outs_dev[self], ldo, w_scale};
    rc = gemm_impl(c, K, T, &q, 1, nullptr, 0, 0, VLUT_ALGO_TCP);
    c->push_flags = nullptr;
    if (rc) return rc;
    float *dst[7];
    int nd = 0;
    for (int i = 0; i < n_out; i++)
        if (i != self) dst[nd++] = outs_dev[i];
    cudaError_t e = vlut::launch_push(outs_dev[self], dst, nd, ldo, N, T, NT, TT, ntiles, (const uint32_t *) c->tile_done.p,
                                      c->act.epoch, c->sm_count, c->push_stream);
    if (e != cudaSuccess) return cuda_fail(e, "push launch");
    c->launches++;
    CK(cudaEventRecord(c->push_join, c->push_stream));
    CK(cudaStreamWaitEvent(c->stream, c->push_join, 0));
    c->inflight_unknown = true;
    return VLUT_OK;
}

int vlut_barrier(vlut_ctx *c, void *const *flags_dev, int n_ranks, int rank) {
    if (!c || !flags_dev || n_ranks < 1 || n_ranks > 8 || rank < 0 || rank >= n_ranks) return VLUT_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    uint32_t *f[8];
    for (int i = 0; i < n_ranks; i++) {
        if (!flags_dev[i]) return VLUT_ERR_INVALID;
        f[i] = (uint32_t *) flags_dev[i];
    }
    cudaError_t e = launch_barrier(f, n_ranks, rank, c->stream);
    if (e != cudaSuccess) return cuda_fail(e, "barrier launch");
    c->launches++;
    c->inflight.clear();   // plain launch: everything before it has completed when it runs
    c->inflight_unknown = false;
    c->prev_tracked = false;
    return VLUT_OK;
}

int vlut_mul_mat(vlut_ctx *c, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev, const float *x_dev,
                 int64_t ldx, float *out_dev, int64_t ldo, float w_scale, int algo) {
    if (use_gemv(&algo, type, K, N, T, x_dev, ldx)) {
        const GemmCall q{type, N, t2_dev, out_dev, ldo, w_scale};
        return gemv_impl(c, K, T, &q, 1, x_dev, ldx);
    }
    int rc = quantize_impl(c, type, x_dev, ldx, K, T, algo, N);
    if (rc) return rc;
    return vlut_gemm(c, type, K, N, T, t2_dev, out_dev, ldo, w_scale, algo);
}

int vlut_ctx_set_option(vlut_ctx *c, int option, int value) {
    if (!c) return VLUT_ERR_INVALID;
    switch (option) {
        case VLUT_OPT_INPUTS_VISIBLE: c->inputs_visible = value != 0; return VLUT_OK;
        default: return VLUT_ERR_INVALID;
    }
}

int vlut_scale_rows(vlut_ctx *c, const float *x_dev, int64_t ldx, float *out_dev, int64_t ldo, int64_t N, int64_t T,
                    float scale) {
    if (!c || !x_dev || !out_dev || N <= 0 || T <= 0 || ldx < N || ldo < N) return VLUT_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    note_foreign_write(c);  // this kernel's output may be the next quantise's input
    cudaError_t e = launch_scale_rows(x_dev, ldx, out_dev, ldo, N, T, scale, c->stream);
    if (e != cudaSuccess) return cuda_fail(e, "scale_rows launch");
    c->launches += (T + 65534) / 65535;
    return VLUT_OK;
}

// rows x width bytes between host and device; contiguous on both sides -> one 1-D copy (the 2-D path is slower per byte)
static cudaError_t copy_rows(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t rows,
                             cudaMemcpyKind kind, cudaStream_t st) {
    if (dpitch == width && spitch == width) return cudaMemcpyAsync(dst, src, width * rows, kind, st);
    return cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, kind, st);
}

// Token chunk of the host-buffer pipeline.  Quantisation is per token, so cutting T is exact; 128 tokens is one TCP token
// tile (GEMM ~11 us) against ~130 us of D2H for the chunk's 7.3 MB of f32 results at 4096->14336.
static int host_chunk_tokens() { return knobs().host_chunk; }

int vlut_mul_mat_host(vlut_ctx *c, int type, int64_t K, int64_t N, int64_t T, const void *t2_dev, const float *x_host,
                      int64_t ldx, float *out_host, int64_t ldo, float w_scale, int algo) {
    if (!c || !x_host || !out_host || T <= 0 || ldx < K || ldo < N) return VLUT_ERR_INVALID;
    int rc = check_shape(type, K, N);
    if (rc) return rc;
    CK(cudaSetDevice(c->device));
    if (c->stage_x.reserve((size_t) (T * K) * sizeof(float)) != cudaSuccess) return VLUT_ERR_NOMEM;
    if (c->stage_out.reserve((size_t) (T * N) * sizeof(float)) != cudaSuccess) return VLUT_ERR_NOMEM;
    float *sx = (float *) c->stage_x.p, *so = (float *) c->stage_out.p;
    const int64_t chunk = host_chunk_tokens();
    if (chunk == 0 || T <= chunk) {
        CK(copy_rows(sx, (size_t) K * 4, x_host, (size_t) ldx * 4, (size_t) K * 4, (size_t) T, cudaMemcpyHostToDevice,
                             c->stream));
        rc = vlut_mul_mat(c, type, K, N, T, t2_dev, sx, K, so, N, w_scale, algo);
        if (rc) return rc;
        CK(copy_rows(out_host, (size_t) ldo * 4, so, (size_t) N * 4, (size_t) N * 4, (size_t) T, cudaMemcpyDeviceToHost,
                             c->stream));
        CK(cudaStreamSynchronize(c->stream));
        return VLUT_OK;
    }
    // Three-stage pipeline over token chunks: H2D of chunk i+1 and D2H of chunk i-1 (PCIe is full duplex: one copy
    // stream per direction) run under the quantise + GEMM of chunk i on the context's stream.  The result copy is the
    // long pole (4TN bytes against 4TK in), so the op costs ~ D2H(out) + one chunk of H2D + one chunk of compute.
    if (!c->h2d_stream) CK(cudaStreamCreateWithFlags(&c->h2d_stream, cudaStreamNonBlocking));
    if (!c->d2h_stream) CK(cudaStreamCreateWithFlags(&c->d2h_stream, cudaStreamNonBlocking));
    // chunk boundaries: the result copy is the long pole and cannot start before the first chunk has been copied in and
    // multiplied, so the first chunk is a quarter of the others (32 + 96 + 128 + 128 + ... tokens)
    std::vector<int64_t> cut;
    cut.push_back(0);
    if (chunk >= 8 && T > chunk) {
        cut.push_back(chunk / 4);
        cut.push_back(chunk);
    }
    while (cut.back() < T) cut.push_back(std::min(T, cut.back() + chunk));
    const int64_t nchunk = (int64_t) cut.size() - 1;
    while ((int64_t) c->ev_in.size() < nchunk) {
        cudaEvent_t a, b;
        CK(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        c->ev_in.push_back(a);
        c->ev_done.push_back(b);
    }
    // the staging buffers may still be read/written by earlier work queued on the context's stream
    CK(cudaEventRecord(c->ev_done[0], c->stream));
    CK(cudaStreamWaitEvent(c->h2d_stream, c->ev_done[0], 0));
    for (int64_t i = 0; i < nchunk; i++) {
        const int64_t t0 = cut[i], tn = cut[i + 1] - cut[i];
        CK(copy_rows(sx + t0 * K, (size_t) K * 4, x_host + t0 * ldx, (size_t) ldx * 4, (size_t) K * 4, (size_t) tn,
                             cudaMemcpyHostToDevice, c->h2d_stream));
        CK(cudaEventRecord(c->ev_in[i], c->h2d_stream));
    }
    for (int64_t i = 0; i < nchunk; i++) {
        const int64_t t0 = cut[i], tn = cut[i + 1] - cut[i];
        CK(cudaStreamWaitEvent(c->stream, c->ev_in[i], 0));
        rc = vlut_mul_mat(c, type, K, N, tn, t2_dev, sx + t0 * K, K, so + t0 * N, N, w_scale, algo);
        if (rc) {
            cudaStreamSynchronize(c->h2d_stream);
            cudaStreamSynchronize(c->stream);
            cudaStreamSynchronize(c->d2h_stream);
            return rc;
        }
        CK(cudaEventRecord(c->ev_done[i], c->stream));
        CK(cudaStreamWaitEvent(c->d2h_stream, c->ev_done[i], 0));
        CK(copy_rows(out_host + t0 * ldo, (size_t) ldo * 4, so + t0 * N, (size_t) N * 4, (size_t) N * 4, (size_t) tn,
                             cudaMemcpyDeviceToHost, c->d2h_stream));
    }
    CK(cudaStreamSynchronize(c->d2h_stream));
    CK(cudaStreamSynchronize(c->stream));
    return VLUT_OK;
}

int vlut_debug_get_act(vlut_ctx *c, int64_t K, int64_t T, int8_t *q_host, float *scales_host) {
    if (!c || c->act_K != K || c->act_T != T) return VLUT_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    if (q_host) {
        int rc = ensure_q_rm(c);
        if (rc) return rc;
    }
    CK(cudaStreamSynchronize(c->stream));
    if (q_host)
        CK(cudaMemcpy2D(q_host, (size_t) K, c->act.q_rm, (size_t) c->act.Kpad, (size_t) K, (size_t) T, cudaMemcpyDeviceToHost));
    if (scales_host) CK(cudaMemcpy(scales_host, c->act.scales, (size_t) T * sizeof(float), cudaMemcpyDeviceToHost));
    return VLUT_OK;
}

int vlut_debug_unpack_t2(vlut_ctx *c, int type, int64_t K, int64_t N, const void *t2_dev, int8_t *w_host) {
    if (!c || !t2_dev || !w_host) return VLUT_ERR_INVALID;
    int rc = check_shape(type, K, N);
    if (rc) return rc;
    CK(cudaSetDevice(c->device));
    int8_t *tmp = nullptr;
    CK(cudaMalloc(&tmp, (size_t) (N * K)));
    cudaError_t e = launch_unpack_t2(K, N, (const uint8_t *) t2_dev, tmp, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = cudaMemcpy(w_host, tmp, (size_t) (N * K), cudaMemcpyDeviceToHost);
    cudaFree(tmp);
    if (e != cudaSuccess) return cuda_fail(e, "unpack_t2");
    c->launches++;
    note_foreign_write(c);
    return VLUT_OK;
}

int vlut_measure_int8_peak(vlut_ctx *c, double *tops, double *clk_per_mma) {
    if (!c || !tops) return VLUT_ERR_INVALID;
    CK(cudaSetDevice(c->device));
    CK(cudaStreamSynchronize(c->stream));
    cudaError_t e = measure_int8_peak(c->sm_count, c->stream, 8192, tops, clk_per_mma);
    if (e != cudaSuccess) return cuda_fail(e, "measure_int8_peak");
    c->launches += 4;
    note_foreign_write(c);
    return VLUT_OK;
}

int vlut_debug_tc_counters(long long *out_host, int n_words) {
    if (!out_host || n_words <= 0 || n_words > vlut::DBG_WORDS) return VLUT_ERR_INVALID;
    if (!vlut::g_dbg_buf) return VLUT_ERR_UNSUPPORTED;
    if (cudaDeviceSynchronize() != cudaSuccess) return VLUT_ERR_CUDA;
    if (cudaMemcpy(out_host, vlut::g_dbg_buf, (size_t) n_words * sizeof(long long), cudaMemcpyDeviceToHost) != cudaSuccess)
        return VLUT_ERR_CUDA;
    return VLUT_OK;
}

}  // extern "C"
