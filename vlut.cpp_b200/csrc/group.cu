// group.cu -- vlut_group: ONE process driving G GPUs of one box for output-row tensor parallelism, behind the C boundary.
//
// The reference splits the rows of one MUL_MAT over CPU threads (ggml/src/ggml-cpu/ggml-cpu.c:1413-1419: thread ith computes
// rows [ith*N/nth, (ith+1)*N/nth)); a group splits them over GPUs.  Device 0 of the group is the ROOT: the activations live
// there and the gathered result lands there.  Every member
//   * holds its row shard of each weight (cut out of the GGUF packing on the host at load: vlut_shard_rows_host),
//   * quantises the activations itself, reading x from the root over NVLink (CUDA peer access),
//   * computes its [T][rows] slice and its GEMM EPILOGUE stores the slice straight into the destination buffer(s) on the
//     other GPU(s) -- peer stores, 16 bytes per lane -- so the gather is fused into the GEMM (vlut_mul_mat_scatter).
// Ordering is by CUDA events between the members' streams (one process: no device-side barrier needed): members wait for
// the root's stream before reading x, the root's stream waits for every member before anything reads the result.
// Hosts: the ggml plugin with VLUT_B200_GPUS=G (one logical ggml device that owns G GPUs), and any FFI
// caller.  The one-process-per-GPU flavour (torch.distributed + symmetric memory + vlut_barrier) is vlut.cpp_b200/shard.py.
#include <new>
#include <vector>

#include "vlut_common.cuh"

using namespace vlut;

struct vlut_group {
    int n = 0;
    std::vector<int> dev;
    std::vector<vlut_ctx *> ctx;
    std::vector<cudaEvent_t> done;
    cudaEvent_t ready = nullptr;
};

struct vlut_group_weight {
    int type = 0;
    int64_t K = 0, N = 0;
    std::vector<int64_t> row0, rows;   // per member
    std::vector<void *> t2;            // per member, device memory on that member's GPU
    bool member0_external = false;     // member 0's shard lives in caller-owned memory (the ggml buffer)
};

extern "C" {

int vlut_group_create(int n_dev, const int *devices, vlut_group **out) {
    if (!out || !devices || n_dev < 1 || n_dev > 8) return VLUT_ERR_INVALID;
    *out = nullptr;
    vlut_group *g = new (std::nothrow) vlut_group();
    if (!g) return VLUT_ERR_NOMEM;
    g->n = n_dev;
    for (int i = 0; i < n_dev; i++) {
        vlut_ctx *c = nullptr;
        int rc = vlut_ctx_create(devices[i], nullptr, &c);
        if (rc != VLUT_OK) {
            vlut_group_destroy(g);
            return rc;
        }
        g->dev.push_back(devices[i]);
        g->ctx.push_back(c);
    }
    // every member reads x from, and stores results into, every other member's memory
    for (int i = 0; i < n_dev; i++) {
        cudaSetDevice(devices[i]);
        for (int j = 0; j < n_dev; j++) {
            if (devices[i] == devices[j]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[i], devices[j]);
            if (!can) {
                vlut_group_destroy(g);
                return VLUT_ERR_UNSUPPORTED;
            }
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                vlut_group_destroy(g);
                return VLUT_ERR_CUDA;
            }
            cudaGetLastError();
        }
        cudaEvent_t ev;
        if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) {
            vlut_group_destroy(g);
            return VLUT_ERR_CUDA;
        }
        g->done.push_back(ev);
    }
    cudaSetDevice(devices[0]);
    if (cudaEventCreateWithFlags(&g->ready, cudaEventDisableTiming) != cudaSuccess) {
        vlut_group_destroy(g);
        return VLUT_ERR_CUDA;
    }
    *out = g;
    return VLUT_OK;
}

void vlut_group_destroy(vlut_group *g) {
    if (!g) return;
    for (size_t i = 0; i < g->ctx.size(); i++) {
        cudaSetDevice(g->dev[i]);
        if (i < g->done.size()) cudaEventDestroy(g->done[i]);
        vlut_ctx_destroy(g->ctx[i]);
    }
    if (g->ready) {
        cudaSetDevice(g->dev[0]);
        cudaEventDestroy(g->ready);
    }
    delete g;
}

int vlut_group_size(const vlut_group *g) { return g ? g->n : 0; }
vlut_ctx *vlut_group_ctx(vlut_group *g, int i) { return (g && i >= 0 && i < g->n) ? g->ctx[i] : nullptr; }

// rows of member i: whole 128-row strips, shards differ by at most one strip (the same plan as shard.py's RowPlan)
int vlut_group_rows(const vlut_group *g, int64_t N, int i, int64_t *row0, int64_t *rows) {
    if (!g || i < 0 || i >= g->n || N <= 0 || !row0 || !rows) return VLUT_ERR_INVALID;
    const int64_t strips = ceil_div(N, TILE_ROWS), base = strips / g->n, extra = strips % g->n;
    int64_t b = 0;
    for (int j = 0; j <= i; j++) {
        const int64_t r = std::min<int64_t>(N - b, (base + (j < extra ? 1 : 0)) * TILE_ROWS);
        if (j == i) {
            *row0 = b;
            *rows = r < 0 ? 0 : r;
        }
        b += r < 0 ? 0 : r;
    }
    return VLUT_OK;
}

size_t vlut_group_shard0_nbytes(const vlut_group *g, int type, int64_t K, int64_t N) {
    int64_t r0 = 0, rows = 0;
    if (vlut_group_rows(g, N, 0, &r0, &rows) != VLUT_OK || rows <= 0) return 0;
    return vlut_repacked_nbytes(type, K, rows);
}

int vlut_group_upload_weights(vlut_group *g, int type, int64_t K, int64_t N, const void *gguf_host, void *shard0_dev,
                              vlut_group_weight **out) {
    if (!g || !gguf_host || !out) return VLUT_ERR_INVALID;
    *out = nullptr;
    if (vlut_gguf_nbytes(type, K, N) == 0) return VLUT_ERR_INVALID;
    vlut_group_weight *w = new (std::nothrow) vlut_group_weight();
    if (!w) return VLUT_ERR_NOMEM;
    w->type = type;
    w->K = K;
    w->N = N;
    w->member0_external = shard0_dev != nullptr;
    std::vector<uint8_t> shard;
    for (int i = 0; i < g->n; i++) {
        int64_t r0 = 0, rows = 0;
        vlut_group_rows(g, N, i, &r0, &rows);
        w->row0.push_back(r0);
        w->rows.push_back(rows);
        w->t2.push_back(nullptr);
        if (rows <= 0) continue;
        void *dst = (i == 0 && shard0_dev) ? shard0_dev : nullptr;
        int rc = VLUT_OK;
        if (!dst) rc = vlut_dev_malloc(g->dev[i], vlut_repacked_nbytes(type, K, rows), &dst);
        if (rc == VLUT_OK) {
            w->t2[i] = dst;
            const void *src = gguf_host;
            if (g->n > 1) {
                shard.resize(vlut_gguf_nbytes(type, K, rows));
                rc = vlut_shard_rows_host(type, K, N, r0, rows, gguf_host, shard.data());
                src = shard.data();
            }
            if (rc == VLUT_OK) rc = vlut_upload_weights(g->ctx[i], type, K, rows, src, dst);
        }
        if (rc != VLUT_OK) {
            vlut_group_free_weights(g, w);
            return rc;
        }
    }
    *out = w;
    return VLUT_OK;
}

void vlut_group_free_weights(vlut_group *g, vlut_group_weight *w) {
    if (!g || !w) return;
    for (int i = 0; i < g->n && i < (int) w->t2.size(); i++)
        if (w->t2[i] && !(i == 0 && w->member0_external)) vlut_dev_free(g->dev[i], w->t2[i]);
    delete w;
}

// outs[j] = base of the [T][ldo] f32 result buffer on destination j (n_dst = 1: gather to one buffer, typically on the root;
// n_dst = G: all-gather into a buffer on every member).  x lives on the root.  On return the ROOT's stream is ordered after
// every member's stores: enqueue consumers there (or vlut_ctx_sync the root).
int vlut_group_mul_mat(vlut_group *g, const vlut_group_weight *w, int64_t T, const float *x_root, int64_t ldx,
                       float *const *outs, int n_dst, int64_t ldo, float w_scale) {
    if (!g || !w || !x_root || !outs || n_dst < 1 || n_dst > 8 || T <= 0 || ldo < w->N) return VLUT_ERR_INVALID;
    cudaStream_t s0 = (cudaStream_t) vlut_ctx_stream(g->ctx[0]);
    cudaError_t e;
    if (g->n > 1) {
        cudaSetDevice(g->dev[0]);
        if ((e = cudaEventRecord(g->ready, s0)) != cudaSuccess) return VLUT_ERR_CUDA;
    }
    int rc = VLUT_OK;
    for (int i = 0; i < g->n && rc == VLUT_OK; i++) {
        if (w->rows[i] <= 0) continue;
        cudaStream_t si = (cudaStream_t) vlut_ctx_stream(g->ctx[i]);
        cudaSetDevice(g->dev[i]);
        if (i > 0 && cudaStreamWaitEvent(si, g->ready, 0) != cudaSuccess) return VLUT_ERR_CUDA;
        float *dst[8];
        for (int j = 0; j < n_dst; j++) dst[j] = outs[j] + w->row0[i];
        rc = vlut_mul_mat_scatter(g->ctx[i], w->type, w->K, w->rows[i], T, w->t2[i], x_root, ldx, dst, n_dst, 0, ldo, w_scale,
                                  VLUT_ALGO_AUTO);
        if (rc == VLUT_OK && i > 0 && cudaEventRecord(g->done[i], si) != cudaSuccess) rc = VLUT_ERR_CUDA;
    }
    cudaSetDevice(g->dev[0]);
    for (int i = 1; i < g->n && rc == VLUT_OK; i++)
        if (w->rows[i] > 0 && cudaStreamWaitEvent(s0, g->done[i], 0) != cudaSuccess) rc = VLUT_ERR_CUDA;
    return rc;
}

int vlut_group_sync(vlut_group *g) {
    if (!g) return VLUT_ERR_INVALID;
    int rc = VLUT_OK;
    for (int i = 0; i < g->n; i++) {
        const int r = vlut_ctx_sync(g->ctx[i]);
        if (r != VLUT_OK) rc = r;
    }
    cudaSetDevice(g->dev[0]);
    return rc;
}

}  // extern "C"
