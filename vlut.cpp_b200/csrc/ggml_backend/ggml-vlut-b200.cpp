// ggml-vlut-b200.cpp -- out-of-tree ggml backend for the Vec-LUT ternary MUL_MAT path on NVIDIA B200 (sm_100a).
//
// Built against the UNMODIFIED ggml headers of spinal-cord/vlut.cpp (ggml/include, ggml/src/ggml-backend-impl.h,
// GGML_BACKEND_API_VERSION 2) into libggml-vlut-b200.so, which exports ggml_backend_init / ggml_backend_score and is
// picked up by ggml_backend_load(), i.e. by every tool of the reference when GGML_BACKEND_PATH points at it
// (ggml/src/ggml-backend-reg.cpp:627-631; llama-bench.cpp:2020; common/arg.cpp:874; tests/test-vlut-gemm.cpp:853).
//
// Scope: supports_op is limited to GGML_OP_MUL_MAT whose src0 is a 2-D tensor of a vlut type
// (I2_V, I2_V_4, I2_V_8, I1_V, I1_V_2) living in this backend's buffer type, with f32 src1 and f32 dst -- the op the
// CPU backend runs in ggml_compute_forward_mul_mat's Vec-LUT branch (ggml/src/ggml-cpu/ggml-cpu.c:1372-1421).
// The buffer type repacks GGUF weights into the HBM-coalesced tiled 2-bit layout at load (set_tensor hook).  All
// arithmetic happens in libvlut_b200.so (include/vlut_b200.h); this file contains no CUDA.
//
// The device reports GGML_BACKEND_DEVICE_TYPE_ACCEL so that libllama offers its buffer type to every weight through
// the CPU buffer-type list (src/llama-model.cpp:329-339) and adds the backend to the scheduler
// (src/llama-context.cpp:198-208) while KV cache and all other ops stay on the CPU backend.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "ggml-backend-impl.h"
#include "ggml-backend.h"
#include "ggml-impl.h"
#include "ggml.h"

#include "../../../include/vlut_b200.h"

namespace {

bool is_vlut_type(enum ggml_type t) {
    return t == GGML_TYPE_I2_V || t == GGML_TYPE_I2_V_4 || t == GGML_TYPE_I2_V_8 || t == GGML_TYPE_I1_V ||
           t == GGML_TYPE_I1_V_2;
}

// ---------------------------------------------------------------------------------------------------------------
// device / registry state (static, never freed: the registry leaks handles on purpose, ggml-backend-reg.cpp:229-236)
// ---------------------------------------------------------------------------------------------------------------
struct device_ctx {
    int device = 0;
    std::string name;
    std::string description;
    vlut_ctx *io_ctx = nullptr;  // context used by buffer uploads / downloads
    std::mutex io_mutex;
    ggml_backend_buffer_type buft{};
    // VLUT_B200_GPUS=G > 1: this ONE logical ggml device owns G GPUs.  Weights are row-sharded over
    // them at load, every vlut MUL_MAT runs on all of them with the gather fused into the GEMM epilogues (vlut_group).
    int n_gpus = 1;
    vlut_group *grp = nullptr;
};

struct buffer_ctx {
    device_ctx *dev = nullptr;
    void *base = nullptr;
    // host copies of one-element f32 tensors written through set_tensor (the `.scale` of a ternary weight): the
    // GEMM epilogue takes the scale by value, so graph_compute needs no device read-back.  Guarded by dev->io_mutex.
    std::unordered_map<const void *, float> scalars;
    // group mode: the row shards of every vlut weight of this buffer, keyed by the tensor's data pointer (shard 0 lives
    // there, in the ggml buffer; the other members' shards are the group's own allocations)
    std::unordered_map<const void *, vlut_group_weight *> shards;
};

// call with dev->io_mutex held: bytes [lo, hi) of the buffer are being rewritten
void forget_scalars(buffer_ctx *bc, const void *lo_, size_t n) {
    const char *lo = (const char *) lo_, *hi = lo + n;
    for (auto it = bc->scalars.begin(); it != bc->scalars.end();) {
        const char *p = (const char *) it->first;
        it = (p >= lo && p < hi) ? bc->scalars.erase(it) : std::next(it);
    }
}

// The activations last quantised into the backend's workspace: consecutive MUL_MATs that share src1 (Q/K/V on one
// attention input, up/gate on one FFN input) quantise once.  Valid inside one graph_compute call only.
struct act_cache {
    const ggml_tensor *src1 = nullptr;
    const char *lo = nullptr, *hi = nullptr;  // byte range of the slice that was quantised
    int64_t K = 0, T = 0;
    bool i1_family = false;  // K-group 640 (I1*) vs 512 (I2*)
};

struct backend_ctx {
    device_ctx *dev = nullptr;
    vlut_ctx *ctx = nullptr;
    std::string name;
    act_cache act;
    // counters behind the "vlut_b200_backend_stats" proc address (tests)
    int64_t n_mul_mat = 0, n_quantize = 0, n_fused_scale = 0, n_scale_rows = 0, n_batched = 0, n_sharded = 0;
    bool owns_ctx = true;   // false in group mode: the backend rides the group's root context
};

bool is_i1_family(enum ggml_type t) { return t == GGML_TYPE_I1_V || t == GGML_TYPE_I1_V_2; }

// GGML_OP_MUL of an f32 matrix by a one-element f32 tensor in our buffer type: the per-tensor `.scale` that follows a
// ternary MUL_MAT (src/models/bitnet.cpp:33-34, src/llama-graph.cpp:780-783).  Returns the scalar operand's index.
int scalar_mul_operand(const ggml_tensor *op) {
    if (op->op != GGML_OP_MUL || op->type != GGML_TYPE_F32) return -1;
    for (int j = 1; j >= 0; j--) {
        const ggml_tensor *s = op->src[j], *x = op->src[1 - j];
        if (!s || !x) continue;
        if (s->type != GGML_TYPE_F32 || ggml_nelements(s) != 1) continue;
        if (x->type != GGML_TYPE_F32 || !ggml_are_same_shape(x, op)) continue;
        if (x->nb[0] != sizeof(float) || op->nb[0] != sizeof(float)) continue;
        return j;
    }
    return -1;
}

// the device's GPU group (group mode only), created on first use
vlut_group *device_group(device_ctx *d) {
    if (d->n_gpus <= 1) return nullptr;
    if (!d->grp) {
        int devs[8];
        // VLUT_B200_GROUP_SAME_DEVICE=1: all members on this one GPU (exercises the group path on a single-GPU box)
        const bool same = getenv("VLUT_B200_GROUP_SAME_DEVICE") != nullptr;
        for (int i = 0; i < d->n_gpus; i++) devs[i] = same ? d->device : d->device + i;
        const int rc = vlut_group_create(d->n_gpus, devs, &d->grp);
        if (rc != VLUT_OK) {
            GGML_LOG_ERROR("%s: vlut_group_create(%d GPUs) failed: %s (%s)\n", __func__, d->n_gpus, vlut_error_string(rc),
                           vlut_last_cuda_error());
            d->grp = nullptr;
        }
    }
    return d->grp;
}

vlut_ctx *io_ctx(device_ctx *d) {
    if (!d->io_ctx) {
        if (vlut_ctx_create(d->device, nullptr, &d->io_ctx) != VLUT_OK) {
            GGML_LOG_ERROR("%s: vlut_ctx_create failed: %s\n", __func__, vlut_last_cuda_error());
            return nullptr;
        }
    }
    return d->io_ctx;
}

// ---------------------------------------------------------------------------------------------------------------
// buffer
// ---------------------------------------------------------------------------------------------------------------
void buffer_free(ggml_backend_buffer_t buffer) {
    buffer_ctx *b = (buffer_ctx *) buffer->context;
    for (auto &kv : b->shards) vlut_group_free_weights(b->dev->grp, kv.second);
    vlut_dev_free(b->dev->device, b->base);
    delete b;
}

void *buffer_get_base(ggml_backend_buffer_t buffer) { return ((buffer_ctx *) buffer->context)->base; }

void buffer_memset_tensor(ggml_backend_buffer_t buffer, ggml_tensor *tensor, uint8_t value, size_t offset, size_t size) {
    buffer_ctx *b = (buffer_ctx *) buffer->context;
    std::lock_guard<std::mutex> lock(b->dev->io_mutex);
    forget_scalars(b, (char *) tensor->data + offset, size);
    vlut_dev_memset(io_ctx(b->dev), (char *) tensor->data + offset, value, size);
}

// The repack hook.  libllama uploads a weight with ONE whole-tensor set_tensor when the device reports no
// async/host_buffer/events caps (src/llama-model-loader.cpp:1003-1007), so the full GGUF-layout tensor is seen here.
void buffer_set_tensor(ggml_backend_buffer_t buffer, ggml_tensor *tensor, const void *data, size_t offset, size_t size) {
    buffer_ctx *b = (buffer_ctx *) buffer->context;
    std::lock_guard<std::mutex> lock(b->dev->io_mutex);
    vlut_ctx *c = io_ctx(b->dev);
    if (is_vlut_type(tensor->type)) {
        GGML_ASSERT(offset == 0 && size == ggml_nbytes(tensor) && "vlut weights must be uploaded whole (they are repacked)");
        GGML_ASSERT(ggml_n_dims(tensor) <= 2);
        forget_scalars(b, tensor->data, vlut_repacked_nbytes((int) tensor->type, tensor->ne[0], tensor->ne[1]));
        if (vlut_group *grp = device_group(b->dev)) {
            // row-shard on the host, upload + repack every member's shard; member 0's goes into the tensor's own storage
            auto old = b->shards.find(tensor->data);
            if (old != b->shards.end()) {
                vlut_group_sync(grp);
                vlut_group_free_weights(grp, old->second);
                b->shards.erase(old);
            }
            vlut_group_weight *gw = nullptr;
            const int rc = vlut_group_upload_weights(grp, (int) tensor->type, tensor->ne[0], tensor->ne[1], data, tensor->data, &gw);
            if (rc != VLUT_OK) GGML_ABORT("vlut_group_upload_weights failed: %s (%s)", vlut_error_string(rc), vlut_last_cuda_error());
            b->shards[tensor->data] = gw;
            return;
        }
        const int rc = vlut_upload_weights(c, (int) tensor->type, tensor->ne[0], tensor->ne[1], data, tensor->data);
        if (rc != VLUT_OK) GGML_ABORT("vlut_upload_weights failed: %s (%s)", vlut_error_string(rc), vlut_last_cuda_error());
        return;
    }
    const int rc = vlut_dev_memcpy_h2d(c, (char *) tensor->data + offset, data, size);
    if (rc != VLUT_OK) GGML_ABORT("H2D copy failed: %s (%s)", vlut_error_string(rc), vlut_last_cuda_error());
    if (tensor->type == GGML_TYPE_F32 && ggml_nelements(tensor) == 1 && offset == 0 && size == sizeof(float)) {
        float v;
        memcpy(&v, data, sizeof(float));
        b->scalars[tensor->data] = v;
    } else {
        forget_scalars(b, (char *) tensor->data + offset, size);
    }
}

void buffer_get_tensor(ggml_backend_buffer_t buffer, const ggml_tensor *tensor, void *data, size_t offset, size_t size) {
    buffer_ctx *b = (buffer_ctx *) buffer->context;
    std::lock_guard<std::mutex> lock(b->dev->io_mutex);
    GGML_ASSERT(!is_vlut_type(tensor->type) && "vlut weights are stored repacked and cannot be read back in GGUF layout");
    const int rc = vlut_dev_memcpy_d2h(io_ctx(b->dev), data, (const char *) tensor->data + offset, size);
    if (rc != VLUT_OK) GGML_ABORT("D2H copy failed: %s (%s)", vlut_error_string(rc), vlut_last_cuda_error());
}

void buffer_clear(ggml_backend_buffer_t buffer, uint8_t value) {
    buffer_ctx *b = (buffer_ctx *) buffer->context;
    std::lock_guard<std::mutex> lock(b->dev->io_mutex);
    b->scalars.clear();
    vlut_dev_memset(io_ctx(b->dev), b->base, value, buffer->size);
}

const ggml_backend_buffer_i buffer_iface = {
    /* .free_buffer   = */ buffer_free,
    /* .get_base      = */ buffer_get_base,
    /* .init_tensor   = */ nullptr,
    /* .memset_tensor = */ buffer_memset_tensor,
    /* .set_tensor    = */ buffer_set_tensor,
    /* .get_tensor    = */ buffer_get_tensor,
    /* .cpy_tensor    = */ nullptr,
    /* .clear         = */ buffer_clear,
    /* .reset         = */ nullptr,
};

// ---------------------------------------------------------------------------------------------------------------
// buffer type
// ---------------------------------------------------------------------------------------------------------------
const char *buft_get_name(ggml_backend_buffer_type_t buft) { return ((device_ctx *) buft->context)->name.c_str(); }

ggml_backend_buffer_t buft_alloc_buffer(ggml_backend_buffer_type_t buft, size_t size) {
    device_ctx *d = (device_ctx *) buft->context;
    void *p = nullptr;
    const int rc = vlut_dev_malloc(d->device, size, &p);
    if (rc != VLUT_OK) {
        GGML_LOG_ERROR("%s: allocating %.2f MiB on %s failed: %s\n", __func__, size / 1024.0 / 1024.0, d->name.c_str(),
                       vlut_error_string(rc));
        return nullptr;  // contract: NULL on OOM
    }
    buffer_ctx *b = new buffer_ctx{d, p};
    return ggml_backend_buffer_init(buft, buffer_iface, b, size);
}

size_t buft_get_alignment(ggml_backend_buffer_type_t) { return 1024; }

// vlut weights are stored in the tiled 2-bit layout: padded to 128x128 tiles, and 2 bits/weight for the 1.6-bit I1 types
size_t buft_get_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor *tensor) {
    const size_t nb = ggml_nbytes(tensor);
    if (is_vlut_type(tensor->type) && ggml_n_dims(tensor) <= 2) {
        const size_t rp = vlut_repacked_nbytes((int) tensor->type, tensor->ne[0], tensor->ne[1]);
        return rp > nb ? rp : nb;
    }
    return nb;
}

bool buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i buft_iface = {
    /* .get_name       = */ buft_get_name,
    /* .alloc_buffer   = */ buft_alloc_buffer,
    /* .get_alignment  = */ buft_get_alignment,
    /* .get_max_size   = */ nullptr,
    /* .get_alloc_size = */ buft_get_alloc_size,
    /* .is_host        = */ buft_is_host,
};

// ---------------------------------------------------------------------------------------------------------------
// backend (stream)
// ---------------------------------------------------------------------------------------------------------------
const char *backend_get_name(ggml_backend_t backend) { return ((backend_ctx *) backend->context)->name.c_str(); }

void backend_free(ggml_backend_t backend) {
    backend_ctx *b = (backend_ctx *) backend->context;
    // VLUT_B200_STATS=1: what ran here, for runs under the reference's own tools (llama-bench has no hook to ask)
    if (getenv("VLUT_B200_STATS"))
        fprintf(stderr, "vlut_b200_backend_stats %s: mul_mat=%lld quantise=%lld fused_scale=%lld scale_rows=%lld batched=%lld "
                "sharded=%lld (x%d GPUs) kernel_launches=%lld\n", b->name.c_str(), (long long) b->n_mul_mat, (long long) b->n_quantize,
                (long long) b->n_fused_scale, (long long) b->n_scale_rows, (long long) b->n_batched, (long long) b->n_sharded,
                b->dev->n_gpus, (long long) vlut_ctx_launch_count(b->ctx));
    if (b->owns_ctx) vlut_ctx_destroy(b->ctx);
    else vlut_group_sync(b->dev->grp);
    delete b;
    delete backend;
}

void backend_synchronize(ggml_backend_t backend) {
    backend_ctx *b = (backend_ctx *) backend->context;
    if (b->dev->grp) vlut_group_sync(b->dev->grp);
    else vlut_ctx_sync(b->ctx);
}

// value of a one-element f32 tensor living in one of our buffers: the host copy kept by set_tensor, else one 4-byte
// read-back (after which it is cached too)
bool scalar_value(backend_ctx *b, const ggml_tensor *s, float *out) {
    const ggml_tensor *base = s->view_src ? s->view_src : s;
    if (!base->buffer || base->buffer->buft != &b->dev->buft) return false;
    buffer_ctx *bc = (buffer_ctx *) base->buffer->context;
    std::lock_guard<std::mutex> lock(b->dev->io_mutex);
    auto it = bc->scalars.find(s->data);
    if (it != bc->scalars.end()) {
        *out = it->second;
        return true;
    }
    if (vlut_ctx_sync(b->ctx) != VLUT_OK) return false;  // the scalar may have been produced on our stream
    if (vlut_dev_memcpy_d2h(io_ctx(b->dev), out, s->data, sizeof(float)) != VLUT_OK) return false;
    bc->scalars[s->data] = *out;
    return true;
}

// a node of ours is about to write [lo, hi): forget cached state that lives there
void note_write(backend_ctx *b, const ggml_tensor *dst) {
    const char *lo = (const char *) dst->data, *hi = lo + ggml_nbytes(dst);
    if (b->act.src1 && lo < b->act.hi && b->act.lo < hi) b->act = act_cache{};
    const ggml_tensor *base = dst->view_src ? dst->view_src : dst;
    if (base->buffer && base->buffer->buft == &b->dev->buft) {
        buffer_ctx *bc = (buffer_ctx *) base->buffer->context;
        std::lock_guard<std::mutex> lock(b->dev->io_mutex);
        forget_scalars(bc, lo, (size_t) (hi - lo));
    }
}

// GGML_OP_MUL_MAT(vlut weight [K,N], f32 [K,T,...]) -> f32 [N,T,...]: the Vec-LUT branch of
// ggml_compute_forward_mul_mat (ggml-cpu.c:1372-1421): quantise activations per token, Vec-LUT GEMM, fp32 epilogue.
// `out_t` is where the result goes (the MUL_MAT node itself, or the trailing MUL node when its scale is folded into
// the epilogue as w_scale: ((float) acc * s_t) * w_scale, the same two roundings as MUL_MAT followed by MUL).
struct mm_item {
    const ggml_tensor *dst;     // the MUL_MAT node
    const ggml_tensor *out_t;   // where its result is stored
    float w_scale;
};

ggml_status mm_fail(backend_ctx *b, const ggml_tensor *dst, int rc) {
    b->act = act_cache{};
    GGML_LOG_ERROR("%s: vlut MUL_MAT(%s) failed: %s (%s)\n", __func__, dst->name, vlut_error_string(rc), vlut_last_cuda_error());
    return rc == VLUT_ERR_NOMEM ? GGML_STATUS_ALLOC_FAILED : GGML_STATUS_FAILED;
}

// A group of MUL_MATs that read the SAME 2-D activation tensor (Q/K/V on the attention input, up/gate on the FFN input:
// consecutive build_lora_mm calls, src/models/bitnet.cpp:28-60, src/llama-graph.cpp:668-671): one quantisation and, in
// the small-token regime, ONE kernel launch over all the weights (vlut_mul_mat_batch).
ggml_status compute_group(backend_ctx *b, const mm_item *items, int n) {
    const ggml_tensor *src1 = items[0].dst->src[1];
    const int64_t K = src1->ne[0], T = src1->ne[1];
    int types[4];
    int64_t Ns[4], ldos[4];
    const void *t2[4];
    float *outs[4];
    float scales[4];
    for (int i = 0; i < n; i++) {
        const ggml_tensor *src0 = items[i].dst->src[0];
        GGML_ASSERT(src0->ne[0] == K && items[i].out_t->nb[0] == sizeof(float));
        types[i] = (int) src0->type;
        Ns[i] = src0->ne[1];
        t2[i] = src0->data;
        outs[i] = (float *) items[i].out_t->data;
        ldos[i] = (int64_t) (items[i].out_t->nb[1] / sizeof(float));
        scales[i] = items[i].w_scale;
    }
    // the input may already be quantised in the workspace (an earlier MUL_MAT of this graph read it)
    const bool reuse = b->act.src1 == src1 && b->act.K == K && b->act.T == T && b->act.i1_family == is_i1_family(items[0].dst->src[0]->type);
    int rc;
    if (reuse) {
        rc = vlut_gemm_batch(b->ctx, n, types, K, Ns, T, t2, outs, ldos, scales, VLUT_ALGO_AUTO);
    } else {
        rc = vlut_mul_mat_batch(b->ctx, n, types, K, Ns, T, t2, (const float *) src1->data,
                                (int64_t) (src1->nb[1] / sizeof(float)), outs, ldos, scales, VLUT_ALGO_AUTO);
        b->n_quantize++;
        b->act = act_cache{};
        if (rc == VLUT_OK && vlut_ctx_last_algo(b->ctx) != VLUT_ALGO_GEMV) {
            b->act.src1 = src1;
            b->act.lo = (const char *) src1->data;
            b->act.hi = (const char *) src1->data + (size_t) (T - 1) * src1->nb[1] + (size_t) K * sizeof(float);
            b->act.K = K;
            b->act.T = T;
            b->act.i1_family = is_i1_family(items[0].dst->src[0]->type);
        }
    }
    if (rc != VLUT_OK) return mm_fail(b, items[0].dst, rc);
    b->n_mul_mat += n;
    b->n_batched += n;
    for (int i = 0; i < n; i++) note_write(b, items[i].out_t);
    return GGML_STATUS_SUCCESS;
}

ggml_status compute_mul_mat(backend_ctx *b, const ggml_tensor *dst, const ggml_tensor *out_t, float w_scale) {
    const ggml_tensor *src0 = dst->src[0];
    const ggml_tensor *src1 = dst->src[1];
    const int64_t K = src0->ne[0], N = src0->ne[1], T = src1->ne[1];
    GGML_ASSERT(src1->ne[0] == K && dst->ne[0] == N && dst->ne[1] == T);
    GGML_ASSERT(src1->nb[0] == sizeof(float) && out_t->nb[0] == sizeof(float));
    const bool batched = src1->ne[2] * src1->ne[3] > 1;
    // group mode: the weight is row-sharded over the device's GPUs
    vlut_group_weight *gw = nullptr;
    if (b->dev->grp) {
        const ggml_tensor *wbase = src0->view_src ? src0->view_src : src0;
        buffer_ctx *bc = wbase->buffer ? (buffer_ctx *) wbase->buffer->context : nullptr;
        std::lock_guard<std::mutex> lock(b->dev->io_mutex);
        auto it = bc ? bc->shards.find(src0->data) : decltype(bc->shards.end()){};
        if (!bc || it == bc->shards.end()) {
            GGML_LOG_ERROR("%s: no row shards for weight %s\n", __func__, src0->name);
            return GGML_STATUS_FAILED;
        }
        gw = it->second;
    }
    for (int64_t i3 = 0; i3 < src1->ne[3]; i3++) {
        for (int64_t i2 = 0; i2 < src1->ne[2]; i2++) {
            const float *x = (const float *) ((const char *) src1->data + i2 * src1->nb[2] + i3 * src1->nb[3]);
            float *out = (float *) ((char *) out_t->data + i2 * out_t->nb[2] + i3 * out_t->nb[3]);
            const int64_t ldx = (int64_t) (src1->nb[1] / sizeof(float)), ldo = (int64_t) (out_t->nb[1] / sizeof(float));
            int rc = VLUT_OK;
            if (gw) {
                // every member quantises x (peer read from the root), computes its rows and stores its slice straight into
                // `out` on the root; on return the root's stream is ordered after all of them
                float *dst1[1] = {out};
                rc = vlut_group_mul_mat(b->dev->grp, gw, T, x, ldx, dst1, 1, ldo, w_scale);
                if (rc != VLUT_OK) return mm_fail(b, dst, rc);
                b->n_mul_mat++;
                b->n_quantize++;
                b->n_sharded++;
                continue;
            }
            const bool reuse = !batched && b->act.src1 == src1 && b->act.K == K && b->act.T == T &&
                               b->act.i1_family == is_i1_family(src0->type);
            if (reuse) {
                rc = vlut_gemm(b->ctx, (int) src0->type, K, N, T, src0->data, out, ldo, w_scale, VLUT_ALGO_AUTO);
            } else {
                // the whole op: the library may run it as ONE kernel at a handful of tokens (quantisation fused), in which
                // case the workspace holds no quantisation to reuse
                rc = vlut_mul_mat(b->ctx, (int) src0->type, K, N, T, src0->data, x, ldx, out, ldo, w_scale, VLUT_ALGO_AUTO);
                b->n_quantize++;
                b->act = act_cache{};
                if (rc == VLUT_OK && !batched && vlut_ctx_last_algo(b->ctx) != VLUT_ALGO_GEMV) {
                    b->act.src1 = src1;
                    b->act.lo = (const char *) x;
                    b->act.hi = (const char *) x + (size_t) (T - 1) * src1->nb[1] + (size_t) K * sizeof(float);
                    b->act.K = K;
                    b->act.T = T;
                    b->act.i1_family = is_i1_family(src0->type);
                }
            }
            if (rc != VLUT_OK) return mm_fail(b, dst, rc);
            b->n_mul_mat++;
        }
    }
    note_write(b, out_t);
    return GGML_STATUS_SUCCESS;
}

// stand-alone MUL by a one-element tensor (the product had other consumers, or did not directly follow its MUL_MAT)
ggml_status compute_scalar_mul(backend_ctx *b, const ggml_tensor *dst) {
    const int j = scalar_mul_operand(dst);
    GGML_ASSERT(j >= 0);
    const ggml_tensor *x = dst->src[1 - j];
    float sc = 1.0f;
    if (!scalar_value(b, dst->src[j], &sc)) return GGML_STATUS_FAILED;
    for (int64_t i3 = 0; i3 < dst->ne[3]; i3++) {
        for (int64_t i2 = 0; i2 < dst->ne[2]; i2++) {
            const float *xi = (const float *) ((const char *) x->data + i2 * x->nb[2] + i3 * x->nb[3]);
            float *oi = (float *) ((char *) dst->data + i2 * dst->nb[2] + i3 * dst->nb[3]);
            const int rc = vlut_scale_rows(b->ctx, xi, (int64_t) (x->nb[1] / sizeof(float)), oi,
                                           (int64_t) (dst->nb[1] / sizeof(float)), dst->ne[0], dst->ne[1], sc);
            if (rc != VLUT_OK) {
                GGML_LOG_ERROR("%s: vlut_scale_rows(%s) failed: %s\n", __func__, dst->name, vlut_error_string(rc));
                return GGML_STATUS_FAILED;
            }
        }
    }
    b->n_scale_rows++;
    note_write(b, dst);
    return GGML_STATUS_SUCCESS;
}

// MUL_MAT node i, with its trailing MUL by the weight's {1} scale folded in when ggml_can_fuse allows (same shape, single
// use, not a graph output: ggml/src/ggml-impl.h:604-643).  Returns how many graph nodes the item covers (1 or 2).
int make_item(backend_ctx *b, ggml_cgraph *cgraph, int i, bool no_fuse, mm_item *it) {
    static const enum ggml_op fuse_ops[2] = {GGML_OP_MUL_MAT, GGML_OP_MUL};
    ggml_tensor *node = cgraph->nodes[i];
    *it = mm_item{node, node, 1.0f};
    float sc = 1.0f;
    if (!no_fuse && ggml_can_fuse(cgraph, i, fuse_ops, 2)) {
        const ggml_tensor *mul = cgraph->nodes[i + 1];
        const int j = scalar_mul_operand(mul);
        if (j >= 0 && mul->src[1 - j] == node && scalar_value(b, mul->src[j], &sc)) {
            it->out_t = mul;
            it->w_scale = sc;
            return 2;
        }
    }
    return 1;
}

bool ranges_overlap(const ggml_tensor *a, const ggml_tensor *c) {
    const char *al = (const char *) a->data, *ah = al + ggml_nbytes(a);
    const char *cl = (const char *) c->data, *ch = cl + ggml_nbytes(c);
    return al < ch && cl < ah;
}

ggml_status backend_graph_compute(ggml_backend_t backend, ggml_cgraph *cgraph) {
    backend_ctx *b = (backend_ctx *) backend->context;
    static const bool no_fuse = getenv("VLUT_B200_NO_FUSE") != nullptr;
    static const bool no_batch = getenv("VLUT_B200_NO_BATCH") != nullptr;
    b->act = act_cache{};  // inputs may have been rewritten since the last graph
    for (int i = 0; i < cgraph->n_nodes; i++) {
        ggml_tensor *node = cgraph->nodes[i];
        // empty tensors (a ubatch without output tokens leaves the last layer's FFN with zero rows) are no-ops, as in
        // every other backend
        if (ggml_is_empty(node)) continue;
        switch (node->op) {
            case GGML_OP_MUL_MAT: {
                mm_item items[4];
                int n = 0, next = i;
                next += make_item(b, cgraph, i, no_fuse, &items[n++]);
                // look ahead: the following MUL_MATs that read the same (2-D) activation tensor with the same K-group
                // family join the group, as long as no member's output overlaps the shared input or another output
                const ggml_tensor *src1 = node->src[1];
                const bool groupable = !no_batch && !b->dev->grp && src1->ne[2] * src1->ne[3] == 1;
                while (groupable && n < 4 && next < cgraph->n_nodes) {
                    ggml_tensor *cand = cgraph->nodes[next];
                    if (cand->op != GGML_OP_MUL_MAT || cand->src[1] != src1 || !cand->src[0] || ggml_is_empty(cand) ||
                        !ggml_backend_dev_supports_op(backend->device, cand) ||
                        is_i1_family(cand->src[0]->type) != is_i1_family(node->src[0]->type))
                        break;
                    mm_item it;
                    const int adv = make_item(b, cgraph, next, no_fuse, &it);
                    bool clash = ranges_overlap(it.out_t, src1);
                    for (int k = 0; k < n; k++) clash = clash || ranges_overlap(it.out_t, items[k].out_t);
                    if (clash) break;
                    items[n++] = it;
                    next += adv;
                }
                if (ranges_overlap(items[0].out_t, src1)) n = 1;
                ggml_status st;
                if (n > 1) {
                    st = compute_group(b, items, n);
                } else {
                    st = compute_mul_mat(b, items[0].dst, items[0].out_t, items[0].w_scale);
                    next = i + (items[0].out_t != items[0].dst ? 2 : 1);
                }
                if (st != GGML_STATUS_SUCCESS) return st;
                for (int k = 0; k < n; k++) b->n_fused_scale += (items[k].out_t != items[k].dst) ? 1 : 0;
                i = next - 1;
            } break;
            case GGML_OP_MUL: {
                const ggml_status st = compute_scalar_mul(b, node);
                if (st != GGML_STATUS_SUCCESS) return st;
            } break;
            case GGML_OP_NONE:
            case GGML_OP_RESHAPE:
            case GGML_OP_VIEW:
            case GGML_OP_PERMUTE:
            case GGML_OP_TRANSPOSE:
                break;
            default:
                GGML_LOG_ERROR("%s: unsupported op %s\n", __func__, ggml_op_desc(node));
                return GGML_STATUS_FAILED;
        }
    }
    b->act = act_cache{};
    // results are read back through get_tensor (which synchronises its own stream): make them visible first
    if (vlut_ctx_sync(b->ctx) != VLUT_OK) return GGML_STATUS_FAILED;
    return GGML_STATUS_SUCCESS;
}

const ggml_backend_i backend_iface = {
    /* .get_name           = */ backend_get_name,
    /* .free               = */ backend_free,
    /* .set_tensor_async   = */ nullptr,
    /* .get_tensor_async   = */ nullptr,
    /* .cpy_tensor_async   = */ nullptr,
    /* .synchronize        = */ backend_synchronize,
    /* .graph_plan_create  = */ nullptr,
    /* .graph_plan_free    = */ nullptr,
    /* .graph_plan_update  = */ nullptr,
    /* .graph_plan_compute = */ nullptr,
    /* .graph_compute      = */ backend_graph_compute,
    /* .event_record       = */ nullptr,
    /* .event_wait         = */ nullptr,
    /* .graph_optimize     = */ nullptr,
};

ggml_guid_t backend_guid() {
    static ggml_guid guid = {0x76, 0x6c, 0x75, 0x74, 0x2d, 0x62, 0x32, 0x30, 0x30, 0x2d, 0x73, 0x6d, 0x31, 0x30, 0x30, 0x61};
    return &guid;
}

// ---------------------------------------------------------------------------------------------------------------
// device
// ---------------------------------------------------------------------------------------------------------------
const char *dev_get_name(ggml_backend_dev_t dev) { return ((device_ctx *) dev->context)->name.c_str(); }
const char *dev_get_description(ggml_backend_dev_t dev) { return ((device_ctx *) dev->context)->description.c_str(); }

void dev_get_memory(ggml_backend_dev_t dev, size_t *free, size_t *total) {
    device_ctx *d = (device_ctx *) dev->context;
    *free = *total = 0;
    vlut_device_info(d->device, nullptr, 0, free, total, nullptr);
}

enum ggml_backend_dev_type dev_get_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_ACCEL; }

void dev_get_props(ggml_backend_dev_t dev, ggml_backend_dev_props *props) {
    props->name = dev_get_name(dev);
    props->description = dev_get_description(dev);
    props->type = dev_get_type(dev);
    props->device_id = nullptr;
    dev_get_memory(dev, &props->memory_free, &props->memory_total);
    // no async / host buffer / events: makes libllama upload every weight with one whole-tensor set_tensor (repack hook)
    props->caps = {/* .async = */ false, /* .host_buffer = */ false, /* .buffer_from_host_ptr = */ false, /* .events = */ false};
}

ggml_backend_t dev_init_backend(ggml_backend_dev_t dev, const char * /*params*/) {
    device_ctx *d = (device_ctx *) dev->context;
    backend_ctx *b = new backend_ctx;
    b->dev = d;
    b->name = d->name;
    if (vlut_group *grp = device_group(d)) {
        // group mode: the backend's stream is the group's root stream (member streams are ordered against it by events)
        b->ctx = vlut_group_ctx(grp, 0);
        b->owns_ctx = false;
        return new ggml_backend{
            /* .guid    = */ backend_guid(),
            /* .iface   = */ backend_iface,
            /* .device  = */ dev,
            /* .context = */ b,
        };
    }
    if (d->n_gpus > 1) {
        delete b;
        return nullptr;
    }
    const int rc = vlut_ctx_create(d->device, nullptr, &b->ctx);
    if (rc != VLUT_OK) {
        GGML_LOG_ERROR("%s: vlut_ctx_create(%d) failed: %s (%s)\n", __func__, d->device, vlut_error_string(rc), vlut_last_cuda_error());
        delete b;
        return nullptr;
    }
    // src1 reaches this backend through a synchronous set_tensor or from the backend's own kernels (no cpy_tensor, no
    // foreign kernels on the context's private stream): the quantise kernel may read it ahead of its dependency wait
    vlut_ctx_set_option(b->ctx, VLUT_OPT_INPUTS_VISIBLE, 1);
    return new ggml_backend{
        /* .guid    = */ backend_guid(),
        /* .iface   = */ backend_iface,
        /* .device  = */ dev,
        /* .context = */ b,
    };
}

ggml_backend_buffer_type_t dev_get_buffer_type(ggml_backend_dev_t dev) { return &((device_ctx *) dev->context)->buft; }

// libllama probes with a dummy MUL_MAT(w, f32[K,512]) whose w->buffer is a 0-byte buffer of the candidate buft
// (src/llama-model.cpp:172-306): decide from the op's types and from where src0 lives.
bool dev_supports_op(ggml_backend_dev_t dev, const ggml_tensor *op) {
    device_ctx *d = (device_ctx *) dev->context;
    switch (op->op) {
        case GGML_OP_NONE:
        case GGML_OP_RESHAPE:
        case GGML_OP_VIEW:
        case GGML_OP_PERMUTE:
        case GGML_OP_TRANSPOSE:
            return true;
        case GGML_OP_MUL_MAT: {
            const ggml_tensor *src0 = op->src[0];
            const ggml_tensor *src1 = op->src[1];
            if (!src0 || !src1) return false;
            if (!is_vlut_type(src0->type)) return false;
            if (src0->ne[2] != 1 || src0->ne[3] != 1) return false;
            if (src1->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32) return false;
            if (src1->nb[0] != sizeof(float) || src1->nb[1] % sizeof(float) != 0) return false;
            // exactly what the launch path accepts (packing constraints, at most 64 K-groups)
            if (!vlut_op_supported((int) src0->type, src0->ne[0], src0->ne[1], src1->ne[1] > 0 ? src1->ne[1] : 1)) return false;
            // the weight must live (or be about to live) in our buffer type: it is stored repacked there
            if (src0->buffer && src0->buffer->buft != &d->buft) return false;
            return true;
        }
        case GGML_OP_MUL: {
            // only the per-tensor `.scale` multiply: f32 matrix x one-element f32 tensor that lives in our buffer type
            const int j = scalar_mul_operand(op);
            if (j < 0) return false;
            const ggml_tensor *sc = op->src[j]->view_src ? op->src[j]->view_src : op->src[j];
            return sc->buffer && sc->buffer->buft == &d->buft;
        }
        default:
            return false;
    }
}

bool dev_supports_buft(ggml_backend_dev_t dev, ggml_backend_buffer_type_t buft) {
    return buft == &((device_ctx *) dev->context)->buft;
}

const ggml_backend_device_i device_iface = {
    /* .get_name             = */ dev_get_name,
    /* .get_description      = */ dev_get_description,
    /* .get_memory           = */ dev_get_memory,
    /* .get_type             = */ dev_get_type,
    /* .get_props            = */ dev_get_props,
    /* .init_backend         = */ dev_init_backend,
    /* .get_buffer_type      = */ dev_get_buffer_type,
    /* .get_host_buffer_type = */ nullptr,
    /* .buffer_from_host_ptr = */ nullptr,
    /* .supports_op          = */ dev_supports_op,
    /* .supports_buft        = */ dev_supports_buft,
    /* .offload_op           = */ nullptr,
    /* .event_new            = */ nullptr,
    /* .event_free           = */ nullptr,
    /* .event_synchronize    = */ nullptr,
};

// ---------------------------------------------------------------------------------------------------------------
// registry
// ---------------------------------------------------------------------------------------------------------------
struct reg_ctx {
    std::vector<device_ctx *> devices;
    std::vector<ggml_backend_device> handles;
};

const char *reg_get_name(ggml_backend_reg_t) { return "VLUT_B200"; }
size_t reg_get_device_count(ggml_backend_reg_t reg) { return ((reg_ctx *) reg->context)->handles.size(); }
ggml_backend_dev_t reg_get_device(ggml_backend_reg_t reg, size_t index) {
    reg_ctx *r = (reg_ctx *) reg->context;
    GGML_ASSERT(index < r->handles.size());
    return &r->handles[index];
}
// "vlut_b200_backend_stats": void(ggml_backend_t, int64_t out[4]) -> {GEMMs, quantise passes, folded scales, stand-alone
// scale kernels} since the backend was created (how tests and llama-bench runs see that the fusions engaged)
void backend_stats(ggml_backend_t backend, int64_t *out) {
    backend_ctx *b = (backend_ctx *) backend->context;
    out[0] = b->n_mul_mat;
    out[1] = b->n_quantize;
    out[2] = b->n_fused_scale;
    out[3] = b->n_scale_rows;
}
// "vlut_b200_backend_stats2": adds out[4] = MUL_MATs that ran inside a batched group, out[5] = MUL_MATs that ran row-sharded
// over the device's GPU group
void backend_stats2(ggml_backend_t backend, int64_t *out) {
    backend_stats(backend, out);
    out[4] = ((backend_ctx *) backend->context)->n_batched;
    out[5] = ((backend_ctx *) backend->context)->n_sharded;
}

void *reg_get_proc_address(ggml_backend_reg_t, const char *name) {
    if (name && strcmp(name, "vlut_b200_backend_stats") == 0) return (void *) backend_stats;
    if (name && strcmp(name, "vlut_b200_backend_stats2") == 0) return (void *) backend_stats2;
    return nullptr;
}

const ggml_backend_reg_i reg_iface = {
    /* .get_name         = */ reg_get_name,
    /* .get_device_count = */ reg_get_device_count,
    /* .get_device       = */ reg_get_device,
    /* .get_proc_address = */ reg_get_proc_address,
};

ggml_backend_reg_t vlut_b200_reg() {
    static ggml_backend_reg reg;
    static bool initialized = false;
    static std::mutex mutex;
    std::lock_guard<std::mutex> lock(mutex);
    if (!initialized) {
        reg_ctx *r = new reg_ctx;
        int n = 0;
        if (vlut_device_count(&n) != VLUT_OK) n = 0;
        // env VLUT_B200_DEVICES limits how many GPUs are exposed as ggml devices
        if (const char *lim = getenv("VLUT_B200_DEVICES")) {
            const int l = atoi(lim);
            if (l >= 0 && l < n) n = l;
        }
        // env VLUT_B200_GPUS=G (2, 4, 8): ONE logical ggml device that owns G GPUs -- weights row-sharded at load, every
        // vlut MUL_MAT runs on all G with the gather fused into the GEMM epilogues (SURVEY.md section 8e "Exposure")
        int gpus = 1;
        if (const char *g = getenv("VLUT_B200_GPUS")) {
            gpus = atoi(g);
            const bool same = getenv("VLUT_B200_GROUP_SAME_DEVICE") != nullptr;
            if (gpus < 1 || gpus > 8 || (!same && gpus > n)) {
                GGML_LOG_ERROR("VLUT_B200_GPUS=%s: %d usable GPU(s); falling back to one GPU per device\n", g, n);
                gpus = 1;
            }
            if (gpus > 1 && n > 1) n = 1;
        }
        reg = ggml_backend_reg{
            /* .api_version = */ GGML_BACKEND_API_VERSION,
            /* .iface       = */ reg_iface,
            /* .context     = */ r,
        };
        r->handles.reserve(n);
        for (int i = 0; i < n; i++) {
            device_ctx *d = new device_ctx;
            d->device = i;
            d->n_gpus = gpus;
            d->name = "VLUT_B200_" + std::to_string(i);
            char gpu[128] = "";
            vlut_device_info(i, gpu, sizeof(gpu), nullptr, nullptr, nullptr);
            d->description = std::string(gpu) + " (Vec-LUT ternary MUL_MAT, sm_100a" +
                             (gpus > 1 ? ", row-sharded over " + std::to_string(gpus) + " GPUs)" : std::string(")"));
            d->buft = ggml_backend_buffer_type{
                /* .iface   = */ buft_iface,
                /* .device  = */ nullptr,  // fixed up below
                /* .context = */ d,
            };
            r->devices.push_back(d);
            r->handles.push_back(ggml_backend_device{
                /* .iface   = */ device_iface,
                /* .reg     = */ &reg,
                /* .context = */ d,
            });
        }
        for (int i = 0; i < n; i++) r->devices[i]->buft.device = &r->handles[i];
        initialized = true;
    }
    return &reg;
}

// 0 = not usable on this system (ggml_backend_load skips it, ggml-backend-reg.cpp:271-300)
int vlut_b200_score() {
    int n = 0;
    if (vlut_device_count(&n) != VLUT_OK) return 0;
    return n > 0 ? 100 : 0;
}

}  // namespace

GGML_BACKEND_DL_IMPL(vlut_b200_reg)
GGML_BACKEND_DL_SCORE_IMPL(vlut_b200_score)
