/*
 * oracle/ref_driver.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Thin driver linked *next to* the reference's own, unmodified kernel sources
 * (compiled where they lie under /root/reference by oracle/Makefile into
 * oracle/_ref/libvlut_ref_<isa>.so).  Nothing in here re-implements arithmetic:
 * it only
 *   - supplies ggml_row_size() for the vlut types (the one libggml-base symbol
 *     ggml/src/ggml-quants-vlut.c needs; semantics of ggml/src/ggml.c:1326-1335),
 *   - lays activations out the way the kernels expect and calls the reference
 *     kernel functions thread-sliced exactly like
 *     ggml/src/ggml-cpu/ggml-cpu.c:1378-1421, except that the activation tile
 *     base uses the CORRECT offset (t/32)*32*K + t%32 (the reference wrapper's
 *     `wdata + i*TABLE_ENTRY_SIZE + j` at ggml-cpu.c:1404 is only right for
 *     T <= 32).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load the resulting library.
 */
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <assert.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef TABLE_ENTRY_SIZE
#define TABLE_ENTRY_SIZE 32
#endif

/* ggml type ids, ggml/include/ggml.h:425-434 */
enum {
    T_I8_V = 39, T_I2_V = 40, T_I2_V_2 = 41, T_I2_V_4 = 42, T_I2_V_8 = 43, T_I2_V_16 = 44,
    T_I1_V = 45, T_I1_V_2 = 46, T_I1_V_4 = 47,
};

/* the only libggml-base symbol the reference packers reference */
size_t ggml_row_size(int type, int64_t ne) {
    if (type == T_I8_V) return (size_t) ne + sizeof(float);
    if (type == T_I1_V || type == T_I1_V_2 || type == T_I1_V_4) {
        assert(ne % 4 == 0);
        return (size_t) (ne / 20 * 4 + ne % 20 / 4);
    }
    assert(ne % 4 == 0);
    return (size_t) (ne / 4);
}

/* reference symbols (ggml/src/ggml-quants-vlut.h, ggml/src/ggml-cpu/ggml-cpu-quants-vlut.h) */
size_t quantize_i2_v  (const float *, void *, int64_t, int64_t, const float *);
size_t quantize_i2_v_4(const float *, void *, int64_t, int64_t, const float *);
size_t quantize_i2_v_8(const float *, void *, int64_t, int64_t, const float *);
size_t quantize_i1_v  (const float *, void *, int64_t, int64_t, const float *);
size_t quantize_i1_v_2(const float *, void *, int64_t, int64_t, const float *);
void quantize_row_i8_v_tile(const float *, void *, int64_t, float *);
typedef void (*gemm_fn)(int, int, int, float *, size_t, const void *, const void *, int, int);
void ggml_gemm_i2v_i8v_lut (int, int, int, float *, size_t, const void *, const void *, int, int);
void ggml_gemm_i2v4_i8v_lut(int, int, int, float *, size_t, const void *, const void *, int, int);
void ggml_gemm_i2v8_i8v_lut(int, int, int, float *, size_t, const void *, const void *, int, int);
void ggml_gemm_i1v_i8v_lut (int, int, int, float *, size_t, const void *, const void *, int, int);
void ggml_gemm_i1v2_i8v_lut(int, int, int, float *, size_t, const void *, const void *, int, int);

static gemm_fn pick_gemm(int type, int *tile) {
    switch (type) { /* type_traits_vlut, ggml-cpu.c:1180-1238 */
        case T_I2_V:   *tile = 1; return ggml_gemm_i2v_i8v_lut;
        case T_I2_V_4: *tile = 4; return ggml_gemm_i2v4_i8v_lut;
        case T_I2_V_8: *tile = 8; return ggml_gemm_i2v8_i8v_lut;
        case T_I1_V:   *tile = 1; return ggml_gemm_i1v_i8v_lut;
        case T_I1_V_2: *tile = 2; return ggml_gemm_i1v2_i8v_lut;
        default: return NULL;
    }
}

size_t vlut_ref_row_size(int type, int64_t K) { return ggml_row_size(type, K); }

/* whole-matrix pack, like llama-quant.cpp:994-1000 -> ggml_quantize_chunk -> quantize_* */
long vlut_ref_pack(int type, const float *w /*[N][K]*/, void *dst, int64_t N, int64_t K) {
    switch (type) {
        case T_I2_V:   return (long) quantize_i2_v  (w, dst, N, K, NULL);
        case T_I2_V_4: return (long) quantize_i2_v_4(w, dst, N, K, NULL);
        case T_I2_V_8: return (long) quantize_i2_v_8(w, dst, N, K, NULL);
        case T_I1_V:   return (long) quantize_i1_v  (w, dst, N, K, NULL);
        case T_I1_V_2: return (long) quantize_i1_v_2(w, dst, N, K, NULL);
        default: return -1;
    }
}

size_t vlut_ref_act_bytes(int64_t K, int64_t T) {
    int64_t Tp = (T + TABLE_ENTRY_SIZE - 1) / TABLE_ENTRY_SIZE * TABLE_ENTRY_SIZE;
    return (size_t) (Tp * K + 4 * T);
}

/* quantise activations into the tile layout (correct base), vy must be vlut_ref_act_bytes(), zeroed here */
void vlut_ref_quantize_act(const float *x, int64_t ldx, int64_t K, int64_t T, void *vy, int nth) {
    const int E = TABLE_ENTRY_SIZE;
    int64_t Tp = (T + E - 1) / E * E;
    memset(vy, 0, (size_t) (Tp * K + 4 * T));
    float *sc = (float *) ((char *) vy + Tp * K);
    if (nth < 1) nth = 1;
#pragma omp parallel for num_threads(nth) schedule(static)
    for (int64_t t = 0; t < T; t++) {
        quantize_row_i8_v_tile(x + t * ldx, (char *) vy + (t / E) * E * K + (t % E), K, sc + t);
    }
}

/* GEMM on pre-quantised activations, thread-sliced over output rows (ggml-cpu.c:1413-1418).
 * I1_V_2's tail addressing needs equal slices (comment at ggml-cpu-quants-vlut.c:1627): the caller
 * must pick nth with N % nth == 0 for that type, or nth = 1. */
int vlut_ref_gemm(int type, int64_t K, int64_t N, int64_t T, const void *w, const void *vy,
                  float *out /*[T][N]*/, int nth) {
    int tile = 0;
    gemm_fn gemm = pick_gemm(type, &tile);
    if (!gemm) return -1;
    if (nth < 1) nth = 1;
#ifndef VLUT_REF_UNEVEN_OK   /* defined for the build with docs/upstream_fixes.patch applied, which lifts the restriction */
    if (type == T_I1_V_2 && K % 20 != 0 && N % nth != 0) return -2;
#endif
#pragma omp parallel for num_threads(nth) schedule(static, 1)
    for (int ith = 0; ith < nth; ith++) {
        int64_t s0 = (ith * N) / nth, s1 = ((ith + 1) * N) / nth;
        if (s0 < s1) {
            gemm(ith, nth, (int) K, out + s0, (size_t) N, (const char *) w + s0 * tile, vy, (int) T, (int) (s1 - s0));
        }
    }
    return 0;
}

/* the full op: f32 activations in, f32 out -- what GGML_OP_MUL_MAT on a vlut weight computes */
int vlut_ref_mul_mat(int type, int64_t K, int64_t N, int64_t T, const void *w, const float *x, int64_t ldx,
                     float *out, int nth) {
    void *vy = malloc(vlut_ref_act_bytes(K, T));
    if (!vy) return -3;
    vlut_ref_quantize_act(x, ldx, K, T, vy, nth);
    int rc = vlut_ref_gemm(type, K, N, T, w, vy, out, nth);
    free(vy);
    return rc;
}

int vlut_ref_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
