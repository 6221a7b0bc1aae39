// llama_logits.cpp -- TEST INFRASTRUCTURE: run one prompt through the reference's UNMODIFIED libllama and dump the logits.
//
// Linked against the reference's own libllama / libggml (built by scripts/build_reference.sh into baseline/_ref/bin).  The
// Vec-LUT B200 plugin is picked up -- or not -- exactly like in every tool of the reference: ggml_backend_load_all()
// honours GGML_BACKEND_PATH (ggml/src/ggml-backend-reg.cpp:627-631).  tests/test_llama_dropin.py runs this binary twice
// (CPU backend only / plugin loaded) and compares the logits.
//
//   vlut_llama_logits <model.gguf> <n_tokens> <n_ubatch> <out.bin> [n_threads]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ggml-backend.h"
#include "llama.h"

int main(int argc, char **argv) {
    if (argc < 5) {
        fprintf(stderr, "usage: %s model.gguf n_tokens n_ubatch out.bin [n_threads]\n", argv[0]);
        return 2;
    }
    const char *model_path = argv[1];
    const int n_tokens = atoi(argv[2]), n_ubatch = atoi(argv[3]);
    const char *out_path = argv[4];
    const int n_threads = argc > 5 ? atoi(argv[5]) : 8;

    ggml_backend_load_all();
    for (size_t i = 0; i < ggml_backend_dev_count(); i++) {
        ggml_backend_dev_t dev = ggml_backend_dev_get(i);
        fprintf(stderr, "device %zu: %s (%s)\n", i, ggml_backend_dev_name(dev), ggml_backend_dev_description(dev));
    }

    llama_model_params mp = llama_model_default_params();
    mp.n_gpu_layers = 0;   // the plugin is an ACCEL device: its buffer type is offered to every weight through the CPU list
    mp.use_mmap = false;
    llama_model *model = llama_model_load_from_file(model_path, mp);
    if (!model) {
        fprintf(stderr, "failed to load %s\n", model_path);
        return 1;
    }
    llama_context_params cp = llama_context_default_params();
    cp.n_ctx = n_tokens + 8;
    cp.n_batch = n_tokens;
    cp.n_ubatch = n_ubatch;
    cp.n_threads = n_threads;
    cp.n_threads_batch = n_threads;
    cp.no_perf = true;
    llama_context *ctx = llama_init_from_model(model, cp);
    if (!ctx) {
        fprintf(stderr, "failed to create the context\n");
        return 1;
    }
    const llama_vocab *vocab = llama_model_get_vocab(model);
    const int n_vocab = llama_vocab_n_tokens(vocab);

    std::vector<llama_token> toks(n_tokens);
    for (int i = 0; i < n_tokens; i++) toks[i] = (llama_token) ((1000003LL * (i + 1) + 12345) % (n_vocab - 300) + 259);
    llama_batch batch = llama_batch_init(n_tokens, 0, 1);
    batch.n_tokens = n_tokens;
    for (int i = 0; i < n_tokens; i++) {
        batch.token[i] = toks[i];
        batch.pos[i] = i;
        batch.n_seq_id[i] = 1;
        batch.seq_id[i][0] = 0;
        batch.logits[i] = 1;
    }
    if (llama_decode(ctx, batch) != 0) {
        fprintf(stderr, "llama_decode failed\n");
        return 1;
    }
    FILE *f = fopen(out_path, "wb");
    if (!f) return 1;
    for (int i = 0; i < n_tokens; i++) {
        const float *lg = llama_get_logits_ith(ctx, i);
        fwrite(lg, sizeof(float), (size_t) n_vocab, f);
    }
    fclose(f);
    fprintf(stderr, "wrote %d x %d logits to %s\n", n_tokens, n_vocab, out_path);
    llama_batch_free(batch);
    llama_free(ctx);
    llama_model_free(model);
    return 0;
}
