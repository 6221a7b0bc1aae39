// cpu_mul_mat.cpp -- TEST INFRASTRUCTURE: one GGML_OP_MUL_MAT on the reference's own CPU path, through the reference's
// ggml API (the call sequence of tests/test-vlut-gemm.cpp: context -> tensors -> graph -> ggml_graph_plan ->
// ggml_graph_compute), linked against libggml-base / libggml-cpu compiled from the reference sources (stock or patched
// with docs/upstream_fixes.patch; scripts/verify_upstream_fixes.sh).
//
//   vlut_cpu_mul_mat <type> <K> <N> <T> <w.bin> <x.bin> <out.bin> <n_threads>
// w.bin = GGUF-packed weight bytes, x.bin = f32 [T][K]; out.bin = f32 [T][N].  Prints the planned work-buffer size and
// what the Vec-LUT path writes into it (ceil32(T) * K bytes of int8 tiles + 4 * T bytes of scales).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ggml-cpu.h"
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
    if (argc < 9) { fprintf(stderr, "usage: see header comment\n"); return 2; }
    const ggml_type type = (ggml_type) atoi(argv[1]);
    const int64_t K = atoll(argv[2]), N = atoll(argv[3]), T = atoll(argv[4]);
    std::vector<char> w = slurp(argv[5]), x = slurp(argv[6]);
    const int n_threads = atoi(argv[8]);

    const size_t need_w = ggml_row_size(type, K) * N;
    if (w.size() != need_w) { fprintf(stderr, "w.bin has %zu bytes, expected %zu\n", w.size(), need_w); return 2; }
    ggml_init_params ip = {/* .mem_size = */ need_w + (size_t) (T * K + T * N) * sizeof(float) + (1u << 20), /* .mem_buffer = */ nullptr,
                           /* .no_alloc = */ false};
    ggml_context *ctx = ggml_init(ip);
    ggml_tensor *a = ggml_new_tensor_2d(ctx, type, K, N);
    ggml_tensor *b = ggml_new_tensor_2d(ctx, GGML_TYPE_F32, K, T);
    memcpy(a->data, w.data(), w.size());
    memcpy(b->data, x.data(), x.size());
    ggml_tensor *c = ggml_mul_mat(ctx, a, b);
    ggml_cgraph *gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, c);
    ggml_cplan plan = ggml_graph_plan(gf, n_threads, nullptr);
    const size_t written = (size_t) ((T + 31) / 32 * 32) * K + 4 * (size_t) T;
    printf("work_size=%zu vlut_path_writes=%zu\n", plan.work_size, written);
    // over-allocate so that an undersized plan does not corrupt the heap of this checker
    std::vector<uint8_t> work(plan.work_size > written ? plan.work_size : written + 64);
    plan.work_data = work.data();
    if (ggml_graph_compute(gf, &plan) != GGML_STATUS_SUCCESS) { fprintf(stderr, "compute failed\n"); return 3; }
    FILE *f = fopen(argv[7], "wb");
    fwrite(c->data, sizeof(float), (size_t) (T * N), f);
    fclose(f);
    ggml_free(ctx);
    return 0;
}
