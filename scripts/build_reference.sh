#!/usr/bin/env bash
# Build the UNMODIFIED reference (spinal-cord/vlut.cpp) tools that drive the hot path -- llama-bench, llama-batched-bench,
# llama-quantize, llama-batched -- plus its ggml libraries into baseline/_ref/bin (git-ignored, shipped to the GPU box).
# Out-of-tree cmake build under /tmp; nothing is written into /root/reference and no reference source is copied.
#   GGML_BACKEND_DL + GGML_CPU_ALL_VARIANTS: the CPU backend is built once per x86 level and picked at run time, so the
#   binaries built in this (GPU-less) container run at full ISA on the GPU box's host; our plugin is loaded through
#   GGML_BACKEND_PATH exactly like any other dynamic backend (ggml/src/ggml-backend-reg.cpp:627-631).
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")/.." && pwd)
BUILD=${BUILD:-/tmp/vlut_refbuild}
OUT=$HERE/baseline/_ref
[ -d "$REF" ] || { echo "reference tree $REF not present: keeping prebuilt $OUT"; exit 0; }
# the environment may export CC/CXX wrappers without OpenMP: use the system compilers
export CC=/usr/bin/gcc CXX=/usr/bin/g++
cmake -S "$REF" -B "$BUILD" -G Ninja -DCMAKE_BUILD_TYPE=Release -DLLAMA_CURL=OFF -DBUILD_SHARED_LIBS=ON \
      -DGGML_BACKEND_DL=ON -DGGML_CPU_ALL_VARIANTS=ON -DGGML_NATIVE=OFF -DLLAMA_BUILD_TESTS=ON -DLLAMA_BUILD_SERVER=OFF \
      -DCMAKE_BUILD_WITH_INSTALL_RPATH=ON -DCMAKE_INSTALL_RPATH='$ORIGIN' > "$BUILD.configure.log" 2>&1
ninja -C "$BUILD" llama-bench llama-batched-bench llama-batched llama-quantize test-vlut-gemm > "$BUILD.build.log" 2>&1
mkdir -p "$OUT/bin" "$OUT/models"
cp -fL "$BUILD"/bin/*.so* "$OUT/bin/"   # SONAME links (libllama.so.0 ...) as plain files
for t in llama-bench llama-batched-bench llama-batched llama-quantize test-vlut-gemm; do cp -f "$BUILD/bin/$t" "$OUT/bin/"; done
# tokenizer metadata for the synthetic model (data, not source)
cp -f "$REF/models/ggml-vocab-llama-spm.gguf" "$OUT/models/"

# test harness on the reference's public API (tests/harness/llama_logits.cpp): logits with / without the plugin
/usr/bin/g++ -O2 -std=c++17 -I"$REF/include" -I"$REF/ggml/include" "$HERE/tests/harness/llama_logits.cpp" \
    -o "$OUT/bin/vlut_llama_logits" -L"$OUT/bin" -lllama -lggml -lggml-base -Wl,-rpath,'$ORIGIN'
