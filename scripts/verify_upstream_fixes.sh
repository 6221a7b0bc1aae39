#!/usr/bin/env bash
# verify_upstream_fixes.sh -- build the reference's ggml CPU libraries twice, stock and with docs/upstream_fixes.patch
# applied to a COPY of the sources under /tmp, plus the one-op checker tests/harness/cpu_mul_mat.cpp against each.
# Output: oracle/_ref/upfix/{stock,patched}/{libggml-base.so,libggml-cpu.so,vlut_cpu_mul_mat} and
# oracle/_ref/upfix/libvlut_ref_patched.so (the two kernel files + oracle/ref_driver.c, like oracle/Makefile's `ref`).
# tests/test_upstream_fixes.py runs the checks.  Needs /root/reference (build container only); nothing is written there.
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
REF="${REF:-/root/reference}"
OUT="$ROOT/oracle/_ref/upfix"
TMP="${TMPDIR:-/tmp}/vlut_upfix_src"
[ -d "$REF/ggml/src" ] || { echo "reference tree $REF not present"; exit 0; }
CC=/usr/bin/gcc; CXX=/usr/bin/g++
[ -x $CC ] || CC=gcc; [ -x $CXX ] || CXX=g++
rm -rf "$TMP"; mkdir -p "$TMP" "$OUT/stock" "$OUT/patched"
cp -r "$REF/ggml" "$TMP/ggml"
(cd "$TMP" && patch -s -p1 < "$ROOT/docs/upstream_fixes.patch")
KNOBS="-DTABLE_ENTRY_SIZE=32 -DWEIGHT_UNROLL_BLOCK=16"
GDEF="-D_GNU_SOURCE -DGGML_VERSION=\"0.9.5\" -DGGML_COMMIT=\"ref\" -DGGML_SCHED_MAX_COPIES=4 -DGGML_BUILD -DGGML_SHARED -DNDEBUG $KNOBS"
BASE_C="ggml.c ggml-alloc.c ggml-quants.c ggml-quants-vlut.c"
BASE_CXX="ggml.cpp ggml-backend.cpp ggml-opt.cpp ggml-threading.cpp gguf.cpp"
CPU_C="ggml-cpu.c quants.c ggml-cpu-quants-vlut.c arch/x86/quants.c"
CPU_CXX="ggml-cpu.cpp repack.cpp hbm.cpp traits.cpp amx/amx.cpp amx/mmq.cpp binary-ops.cpp unary-ops.cpp vec.cpp ops.cpp llamafile/sgemm.cpp arch/x86/repack.cpp arch/x86/cpu-feats.cpp"
build() {   # $1 = ggml source tree, $2 = output dir
  local G="$1" O="$2"; local INC="-I$G/include -I$G/src -I$G/src/ggml-cpu"
  mkdir -p "$O/obj"
  for f in $BASE_C; do $CC -O2 -fPIC -std=gnu11 $GDEF $INC -c "$G/src/$f" -o "$O/obj/base_$(echo $f | tr / _).o" & done
  for f in $BASE_CXX; do $CXX -O2 -fPIC -std=c++17 $GDEF $INC -c "$G/src/$f" -o "$O/obj/base_$(echo $f | tr / _).o" & done
  for f in $CPU_C; do $CC -O2 -march=x86-64-v3 -fPIC -std=gnu11 $GDEF -DGGML_USE_CPU_REPACK -DGGML_USE_LLAMAFILE $INC -c "$G/src/ggml-cpu/$f" -o "$O/obj/cpu_$(echo $f | tr / _).o" & done
  for f in $CPU_CXX; do $CXX -O2 -march=x86-64-v3 -fPIC -std=c++17 $GDEF -DGGML_USE_CPU_REPACK -DGGML_USE_LLAMAFILE $INC -c "$G/src/ggml-cpu/$f" -o "$O/obj/cpu_$(echo $f | tr / _).o" & done
  wait
  $CXX -shared -o "$O/libggml-base.so" "$O"/obj/base_*.o -lm -lpthread
  $CXX -shared -o "$O/libggml-cpu.so" "$O"/obj/cpu_*.o -L"$O" -lggml-base -lm -lpthread -fopenmp -Wl,-rpath,'$ORIGIN'
  $CXX -O2 -std=c++17 $INC -o "$O/vlut_cpu_mul_mat" "$ROOT/tests/harness/cpu_mul_mat.cpp" -L"$O" -lggml-cpu -lggml-base -Wl,-rpath,'$ORIGIN'
  rm -rf "$O/obj"
}
build "$REF/ggml" "$OUT/stock"
build "$TMP/ggml" "$OUT/patched"
G="$TMP/ggml"
$CC -O3 -fPIC -shared -fopenmp -fno-math-errno -Wno-unused-result -march=x86-64-v3 -DVLUT_REF_UNEVEN_OK $KNOBS -I$G/include -I$G/src -I$G/src/ggml-cpu \
    -o "$OUT/libvlut_ref_patched.so" "$G/src/ggml-quants-vlut.c" "$G/src/ggml-cpu/ggml-cpu-quants-vlut.c" "$ROOT/oracle/ref_driver.c" -lm
rm -rf "$TMP"
ls -la "$OUT" "$OUT/stock" "$OUT/patched"
