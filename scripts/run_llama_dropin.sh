#!/usr/bin/env bash
# Drop-in runs of the B200 plugin under the reference's OWN, UNMODIFIED drivers (built by scripts/build_reference.sh into
# baseline/_ref/bin): llama-quantize -> logits check -> llama-bench pp512 -> llama-batched-bench, each once CPU-only and
# once with GGML_BACKEND_PATH pointing at the plugin (tools/llama-bench/llama-bench.cpp:2020, common/arg.cpp:874;
# methodology: evaluation/scripts/bench-prefill.sh:193).  Run on the GPU box: `gpurun -- scripts/run_llama_dropin.sh`.
# Output: gpurun_out/llama_dropin.txt
set -u
ROOT=$(cd "$(dirname "$0")/.." && pwd)
BIN=$ROOT/baseline/_ref/bin
PLUGIN=$ROOT/vlut.cpp_b200/csrc/build/libggml-vlut-b200.so
W=${VLUT_MODEL_DIR:-/tmp/vlut_models}
OUT=$ROOT/gpurun_out/llama_dropin.txt
LAYERS=${VLUT_LLAMA_LAYERS:-32}
SECTIONS=${VLUT_DROPIN_SECTIONS:-"1 2 3"}   # which of the three parts to run
mkdir -p "$W" "$ROOT/gpurun_out"
exec > >(tee "$OUT") 2>&1
NT=$(nproc)
echo "# host: $(grep -m1 'model name' /proc/cpuinfo | cut -d: -f2) x $NT threads; GPU: $(nvidia-smi --query-gpu=name --format=csv,noheader | head -1)"
[ -x "$BIN/llama-bench" ] || { echo "baseline/_ref/bin missing: run scripts/build_reference.sh in the build container"; exit 1; }
FILTER='^llama_model_loader: - kv\|^llama_model_loader: - type\|^print_info\|^load: \|^\.\.\.\|^llama_context:\|^llama_kv_cache\|^graph_reserve\|^create_tensor\|^set_abort'

if [[ " $SECTIONS " == *" 1 "* ]]; then
echo
echo "## 1. stock llama-quantize on a synthetic f32 ternary model, then logits: CPU backend vs plugin (n_ubatch = 32)"
python "$ROOT/scripts/make_synthetic_gguf.py" --out "$W/falcon_f32_2L.gguf" --model falcon3-1b --layers 2 --ftype f32
"$BIN/llama-quantize" "$W/falcon_f32_2L.gguf" "$W/falcon_i2v4_2L.gguf" I2_V_4 2>&1 | grep "converting to\|quant size\|quantize time" | tail -12
rm -f "$W/falcon_f32_2L.gguf"
timeout 300 "$BIN/vlut_llama_logits" "$W/falcon_i2v4_2L.gguf" 64 32 "$W/logits_cpu.bin" "$NT" 2>&1 | grep "^device\|^wrote\|fail"
GGML_BACKEND_PATH=$PLUGIN VLUT_B200_STATS=1 timeout 300 "$BIN/vlut_llama_logits" "$W/falcon_i2v4_2L.gguf" 64 32 "$W/logits_gpu.bin" "$NT" 2>&1 | grep "^device\|^wrote\|fail\|vlut_b200_backend_stats\|VLUT_B200.*buffer size"
python "$ROOT/scripts/compare_logits.py" "$W/logits_cpu.bin" "$W/logits_gpu.bin"
echo "logits check exit code: $?"
fi

if [[ " $SECTIONS " == *" 2 "* ]]; then
echo
echo "## 2. llama-bench -p 512 -n 0: Llama3-8B shapes, I2_V_4, $LAYERS layers, random-init ternary weights"
python "$ROOT/scripts/make_synthetic_gguf.py" --out "$W/llama3_8b_i2v4.gguf" --model llama3-8b --layers "$LAYERS" --ftype i2_v_4
echo "### CPU backend only ($NT threads)"
timeout 900 "$BIN/llama-bench" -m "$W/llama3_8b_i2v4.gguf" -p 512 -n 0 -r 3 -t "$NT" 2>&1 | grep -v "$FILTER" | tail -6
echo "### plugin loaded (GGML_BACKEND_PATH)"
GGML_BACKEND_PATH=$PLUGIN VLUT_B200_STATS=1 timeout 900 "$BIN/llama-bench" -m "$W/llama3_8b_i2v4.gguf" -p 512 -n 0 -r 3 -t "$NT" 2>&1 | grep -v "$FILTER" | tail -8
echo "### plugin, -ub 2048 -p 2048"
GGML_BACKEND_PATH=$PLUGIN timeout 900 "$BIN/llama-bench" -m "$W/llama3_8b_i2v4.gguf" -p 2048 -b 2048 -ub 2048 -n 0 -r 2 -t "$NT" 2>&1 | grep -v "$FILTER" | tail -4
echo "### GGML_SCHED_DEBUG=2 excerpt (which backend runs which node, first graph)"
GGML_BACKEND_PATH=$PLUGIN GGML_SCHED_DEBUG=2 timeout 300 "$BIN/llama-bench" -v -m "$W/llama3_8b_i2v4.gguf" -p 32 -n 0 -r 1 -t "$NT" 2>&1 | grep -m 48 "SPLIT\|MUL_MAT"
rm -f "$W/llama3_8b_i2v4.gguf"
fi

if [[ " $SECTIONS " == *" 3 "* ]]; then
echo
echo "## 3. llama-batched-bench: Falcon3-1B shapes (18 layers), I2_V_4, parallel decode np = 32..256"
python "$ROOT/scripts/make_synthetic_gguf.py" --out "$W/falcon3_1b_i2v4.gguf" --model falcon3-1b --layers 18 --ftype i2_v_4
echo "### CPU backend only ($NT threads)"
timeout 1200 "$BIN/llama-batched-bench" -m "$W/falcon3_1b_i2v4.gguf" -c 16384 -b 2048 -ub 512 -npp 16 -ntg 16 -npl 32,64,128,256 -t "$NT" 2>&1 | grep "^|" | tail -8
echo "### plugin loaded (GGML_BACKEND_PATH)"
GGML_BACKEND_PATH=$PLUGIN VLUT_B200_STATS=1 timeout 1200 "$BIN/llama-batched-bench" -m "$W/falcon3_1b_i2v4.gguf" -c 16384 -b 2048 -ub 512 -npp 16 -ntg 16 -npl 32,64,128,256 -t "$NT" > "$W/bb.log" 2>&1; echo "exit code $?"; grep "^|\|vlut_b200_backend_stats" "$W/bb.log" | tail -9; grep -i "error\|abort\|assert\|fail\|unsupported" "$W/bb.log" | head -10; tail -5 "$W/bb.log"
fi
rm -rf "$W"
echo "# done"
