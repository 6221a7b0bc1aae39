// ggml_harness.cpp -- TEST INFRASTRUCTURE: drives libggml-vlut-b200.so through the reference's own, unmodified ggml
// API (libggml / libggml-base compiled from /root/reference into oracle/_ref by oracle/Makefile), the way the
// reference's tools do: ggml_backend_load -> device -> buffer type -> tensors -> graph -> compute.
//
//   vlut_ggml_harness <plugin.so> <mode> <type> <K> <N> <T> <w.bin> <x.bin> <out.bin> [cpu-backend.so]
//     mode "direct": every tensor lives in the plugin's buffer; ggml_backend_graph_compute on the plugin backend
//                    (the call sequence of tests/test-vlut-gemm.cpp:459-474)
//     mode "sched" : weight in the plugin's buffer, activations and result on the CPU backend; ggml_backend_sched
//                    splits the graph (MUL_MAT -> plugin, following ADD -> CPU) and copies across the boundary, like
//                    llama_decode does (ggml/src/ggml-backend.cpp:1443-1460)
//     mode "block" : like "sched", with the weight's per-tensor scales ({1} f32 tensors, as `.scale` in
//                    src/models/bitnet.cpp:33-34) in the plugin's buffer too.  One input feeds four MUL_MATs:
//                    q = mul(mul_mat(a,b), sa); k = mul(mul_mat(a,b), sb); v = mul_mat(a,b);
//                    g = mul_mat(a,b) with two consumers: r1 = mul(g, sa), r2 = g + 1 (CPU).
//                    out.bin = q,k,v,r1,r2 (5 x [T][N]); prints the plugin's fusion counters.
// w.bin = GGUF-packed weight bytes, x.bin = f32 [T][K]; out.bin = f32 [T][N] (sched mode: mul_mat result + 1.0).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "ggml-alloc.h"
#include "ggml-backend.h"
#include "ggml.h"

static std::vector<char> slurp(const char *path) {
    FILE *f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<char> buf(n);
    if (fread(buf.data(), 1, n, f) != (size_t) n) { fprintf(stderr, "short read %s\n", path); exit(2); }
    fclose(f);
    return buf;
}

int main(int argc, char **argv) {
    if (argc < 10) { fprintf(stderr, "usage: see header comment\n"); return 2; }
    const char *plugin = argv[1];
    const std::string mode = argv[2];
    const ggml_type type = (ggml_type) atoi(argv[3]);
    const int64_t K = atoll(argv[4]), N = atoll(argv[5]), T = atoll(argv[6]);
    std::vector<char> w = slurp(argv[7]), x = slurp(argv[8]);

    ggml_backend_reg_t reg = ggml_backend_load(plugin);
    if (!reg) { fprintf(stderr, "ggml_backend_load(%s) failed\n", plugin); return 3; }
    printf("loaded backend reg '%s' with %zu device(s)\n", ggml_backend_reg_name(reg), ggml_backend_reg_dev_count(reg));
    if (ggml_backend_reg_dev_count(reg) < 1) return 3;
    ggml_backend_dev_t dev = ggml_backend_reg_dev_get(reg, 0);
    size_t mfree = 0, mtotal = 0;
    ggml_backend_dev_memory(dev, &mfree, &mtotal);
    printf("device '%s': %s, type %d, %.1f / %.1f GiB free\n", ggml_backend_dev_name(dev), ggml_backend_dev_description(dev),
           (int) ggml_backend_dev_type(dev), mfree / 1073741824.0, mtotal / 1073741824.0);
    ggml_backend_t backend = ggml_backend_dev_init(dev, nullptr);
    if (!backend) { fprintf(stderr, "backend init failed\n"); return 3; }
    ggml_backend_buffer_type_t buft = ggml_backend_dev_buffer_type(dev);

    ggml_init_params ip = {/* .mem_size = */ ggml_tensor_overhead() * 16 + ggml_graph_overhead(), /* .mem_buffer = */ nullptr,
                           /* .no_alloc = */ true};
    std::vector<float> out((size_t) T * N);

    if (mode == "direct") {
        ggml_context *ctx = ggml_init(ip);
        ggml_tensor *a = ggml_new_tensor_2d(ctx, type, K, N);
        ggml_tensor *b = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, K, T);
        ggml_tensor *c = ggml_mul_mat(ctx, a, b);
        if (w.size() != ggml_nbytes(a) || x.size() != ggml_nbytes(b)) { fprintf(stderr, "size mismatch\n"); return 2; }
        if (!ggml_backend_dev_supports_op(dev, c)) { fprintf(stderr, "supports_op says no\n"); return 4; }
        ggml_backend_buffer_t buf = ggml_backend_alloc_ctx_tensors_from_buft(ctx, buft);
        if (!buf) { fprintf(stderr, "alloc failed\n"); return 5; }
        ggml_backend_tensor_set(a, w.data(), 0, w.size());
        ggml_backend_tensor_set(b, x.data(), 0, x.size());
        ggml_cgraph *gf = ggml_new_graph(ctx);
        ggml_build_forward_expand(gf, c);
        const ggml_status st = ggml_backend_graph_compute(backend, gf);
        if (st != GGML_STATUS_SUCCESS) { fprintf(stderr, "graph_compute status %d\n", (int) st); return 6; }
        ggml_backend_tensor_get(c, out.data(), 0, out.size() * sizeof(float));
        ggml_backend_buffer_free(buf);
        ggml_free(ctx);
    } else if (mode == "sched") {
        if (argc < 11) { fprintf(stderr, "sched mode needs the CPU backend library\n"); return 2; }
        ggml_backend_reg_t cpu_reg = ggml_backend_load(argv[10]);
        if (!cpu_reg) { fprintf(stderr, "loading CPU backend failed\n"); return 3; }
        ggml_backend_dev_t cpu_dev = ggml_backend_reg_dev_get(cpu_reg, 0);
        ggml_backend_t cpu = ggml_backend_dev_init(cpu_dev, nullptr);
        // weights context: in the plugin's buffer type (what libllama's loader does per weight)
        ggml_context *wctx = ggml_init(ip);
        ggml_tensor *a = ggml_new_tensor_2d(wctx, type, K, N);
        ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors_from_buft(wctx, buft);
        if (!wbuf) { fprintf(stderr, "weight alloc failed\n"); return 5; }
        ggml_backend_buffer_set_usage(wbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
        ggml_backend_tensor_set(a, w.data(), 0, w.size());
        // compute graph: c = mul_mat(a, b); d = c + 1 (forces a CPU split after ours)
        ggml_context *gctx = ggml_init(ip);
        ggml_tensor *b = ggml_new_tensor_2d(gctx, GGML_TYPE_F32, K, T);
        ggml_set_input(b);
        ggml_tensor *c = ggml_mul_mat(gctx, a, b);
        ggml_tensor *one = ggml_new_tensor_1d(gctx, GGML_TYPE_F32, 1);
        ggml_set_input(one);
        ggml_tensor *d = ggml_add(gctx, c, one);
        ggml_set_output(d);
        ggml_cgraph *gf = ggml_new_graph(gctx);
        ggml_build_forward_expand(gf, d);
        ggml_backend_t backends[2] = {backend, cpu};
        ggml_backend_sched_t sched = ggml_backend_sched_new(backends, nullptr, 2, 64, false, true);
        if (!ggml_backend_sched_alloc_graph(sched, gf)) { fprintf(stderr, "sched alloc failed\n"); return 5; }
        const float onev = 1.0f;
        ggml_backend_tensor_set(b, x.data(), 0, x.size());
        ggml_backend_tensor_set(one, &onev, 0, sizeof(float));
        const ggml_status st = ggml_backend_sched_graph_compute(sched, gf);
        if (st != GGML_STATUS_SUCCESS) { fprintf(stderr, "sched compute status %d\n", (int) st); return 6; }
        printf("sched: %d splits, mul_mat ran on '%s'\n", ggml_backend_sched_get_n_splits(sched),
               ggml_backend_name(ggml_backend_sched_get_tensor_backend(sched, c)));
        ggml_backend_tensor_get(d, out.data(), 0, out.size() * sizeof(float));
        ggml_backend_sched_free(sched);
        ggml_backend_buffer_free(wbuf);
        ggml_free(gctx);
        ggml_free(wctx);
        ggml_backend_free(cpu);
    } else if (mode == "block") {
        if (argc < 11) { fprintf(stderr, "block mode needs the CPU backend library\n"); return 2; }
        ggml_backend_reg_t cpu_reg = ggml_backend_load(argv[10]);
        if (!cpu_reg) { fprintf(stderr, "loading CPU backend failed\n"); return 3; }
        ggml_backend_t cpu = ggml_backend_dev_init(ggml_backend_reg_dev_get(cpu_reg, 0), nullptr);
        ggml_init_params ipb = {ggml_tensor_overhead() * 64 + ggml_graph_overhead(), nullptr, true};
        ggml_context *wctx = ggml_init(ipb);
        ggml_tensor *a = ggml_new_tensor_2d(wctx, type, K, N);
        ggml_tensor *sa = ggml_new_tensor_1d(wctx, GGML_TYPE_F32, 1);
        ggml_tensor *sb = ggml_new_tensor_1d(wctx, GGML_TYPE_F32, 1);
        ggml_backend_buffer_t wbuf = ggml_backend_alloc_ctx_tensors_from_buft(wctx, buft);
        if (!wbuf) { fprintf(stderr, "weight alloc failed\n"); return 5; }
        ggml_backend_buffer_set_usage(wbuf, GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
        const float sav = 0.37f, sbv = -1.9f, onev = 1.0f;
        ggml_backend_tensor_set(a, w.data(), 0, w.size());
        ggml_backend_tensor_set(sa, &sav, 0, sizeof(float));
        ggml_backend_tensor_set(sb, &sbv, 0, sizeof(float));
        ggml_context *gctx = ggml_init(ipb);
        ggml_tensor *b = ggml_new_tensor_2d(gctx, GGML_TYPE_F32, K, T);
        ggml_set_input(b);
        ggml_tensor *one = ggml_new_tensor_1d(gctx, GGML_TYPE_F32, 1);
        ggml_set_input(one);
        ggml_tensor *q = ggml_mul(gctx, ggml_mul_mat(gctx, a, b), sa);
        ggml_tensor *k = ggml_mul(gctx, ggml_mul_mat(gctx, a, b), sb);
        ggml_tensor *v = ggml_mul_mat(gctx, a, b);
        ggml_tensor *g = ggml_mul_mat(gctx, a, b);
        ggml_tensor *r1 = ggml_mul(gctx, g, sa);
        ggml_tensor *r2 = ggml_add(gctx, g, one);
        ggml_tensor *outs[5] = {q, k, v, r1, r2};
        ggml_cgraph *gf = ggml_new_graph(gctx);
        for (ggml_tensor *o : outs) { ggml_set_output(o); ggml_build_forward_expand(gf, o); }
        ggml_backend_t backends[2] = {backend, cpu};
        ggml_backend_sched_t sched = ggml_backend_sched_new(backends, nullptr, 2, 64, false, true);
        if (!ggml_backend_sched_alloc_graph(sched, gf)) { fprintf(stderr, "sched alloc failed\n"); return 5; }
        ggml_backend_tensor_set(b, x.data(), 0, x.size());
        ggml_backend_tensor_set(one, &onev, 0, sizeof(float));
        const ggml_status st = ggml_backend_sched_graph_compute(sched, gf);
        if (st != GGML_STATUS_SUCCESS) { fprintf(stderr, "sched compute status %d\n", (int) st); return 6; }
        printf("block: %d splits, scaled product ran on '%s'\n", ggml_backend_sched_get_n_splits(sched),
               ggml_backend_name(ggml_backend_sched_get_tensor_backend(sched, q)));
        typedef void (*stats_fn)(ggml_backend_t, int64_t *);
        stats_fn stats = (stats_fn) ggml_backend_reg_get_proc_address(reg, "vlut_b200_backend_stats");
        stats_fn stats2 = (stats_fn) ggml_backend_reg_get_proc_address(reg, "vlut_b200_backend_stats2");
        if (stats) {
            int64_t s4[6] = {0, 0, 0, 0, 0, 0};
            if (stats2) stats2(backend, s4); else stats(backend, s4);
            printf("stats: gemm=%lld quantize=%lld fused_scale=%lld scale_rows=%lld batched=%lld sharded=%lld\n", (long long) s4[0],
                   (long long) s4[1], (long long) s4[2], (long long) s4[3], (long long) s4[4], (long long) s4[5]);
        }
        out.resize((size_t) 5 * T * N);
        for (int i = 0; i < 5; i++)
            ggml_backend_tensor_get(outs[i], out.data() + (size_t) i * T * N, 0, (size_t) T * N * sizeof(float));
        ggml_backend_sched_free(sched);
        ggml_backend_buffer_free(wbuf);
        ggml_free(gctx);
        ggml_free(wctx);
        ggml_backend_free(cpu);
    } else {
        fprintf(stderr, "unknown mode %s\n", mode.c_str());
        return 2;
    }
    ggml_backend_free(backend);
    FILE *f = fopen(argv[9], "wb");
    fwrite(out.data(), sizeof(float), out.size(), f);
    fclose(f);
    printf("ok\n");
    return 0;
}
