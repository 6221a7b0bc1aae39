"""oracle/make_golden.py -- TEST INFRASTRUCTURE: regenerate tests/golden/*.npz.

Run in the build container (needs /root/reference to have been compiled into oracle/_ref by `make -C oracle`):

    python oracle/make_golden.py

Every fixture is an output of the REFERENCE'S OWN compiled kernels (oracle.RefLib): packed weight bytes from
its packers (ggml/src/ggml-quants-vlut.c), int8 activations + scales from quantize_row_i8_v_tile and fp32
results from ggml_gemm_*_i8v_lut (ggml/src/ggml-cpu/ggml-cpu-quants-vlut.c), called with the correct
activation tile layout.  I1 types are only recorded for K % 20 == 0, where the reference
computes its intended math (for K % 20 != 0 its tail table is stale, ggml-cpu-quants-vlut.c:1570-1572).

Also writes tests/golden/kat_full.json: sha256 known answers of full-size configs (BASELINE.json configs[0]
and friends) on seeded inputs, so the GPU path can be checked at sizes whose tensors are too big to commit.
"""
from __future__ import annotations

import hashlib
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import oracle as O  # noqa: E402

OUT = ROOT / "tests" / "golden"

SMALL_CASES = [
    # name,            type,     K,    N,  T,  weights,     acts
    ("i2v_k1024",      O.I2_V,   1024, 48, 33, "ternary",   "normal"),
    ("i2v4_k1280",     O.I2_V_4, 1280, 40, 37, "ternary",   "outlier"),
    ("i2v4_k640_t1",   O.I2_V_4, 640,  64, 1,  "ternary",   "uniform"),
    ("i2v8_k1024",     O.I2_V_8, 1024, 32, 40, "zero_free", "normal"),
    ("i1v_k1280",      O.I1_V,   1280, 48, 35, "ternary",   "normal"),
    ("i1v2_k640",      O.I1_V_2, 640,  64, 7,  "ternary",   "uniform"),
    ("i1v2_k1300",     O.I1_V_2, 1300, 36, 34, "ternary",   "outlier"),   # K%20==0 but K%640!=0: short last group
    ("i2v4_negrows",   O.I2_V_4, 512,  32, 5,  "ternary",   "negative"),  # signed-max rule: everything -> -128
    ("i2v4_wrap",      O.I2_V_4, 2048, 16, 3,  "ones",      "ones"),      # int16 wrap: 4 groups x wrap16(65024)
    ("i1v2_wrap",      O.I1_V_2, 1920, 16, 2,  "ones",      "ones"),      # 3 groups x wrap16(81280)
    ("i2v4_zeros",     O.I2_V_4, 512,  16, 2,  "ternary",   "zeros"),     # all-zero activations: scale floor 1e-5/127
]

# full-size known answers: (name, type, K, N, T)
KAT_CASES = [
    ("cfg0_i2v4_4096x4096x256", O.I2_V_4, 4096, 4096, 256),     # BASELINE.json configs[0]
    ("bitnet_i1v2_3200x8640x65", O.I1_V_2, 3200, 8640, 65),     # configs[1] shape, ragged T
    ("llama_i2v4_4096x14336x16", O.I2_V_4, 4096, 14336, 16),    # headline shape, small-batch regime
]


def gen_weights(kind: str, N: int, K: int, seed: int) -> np.ndarray:
    if kind == "ones":
        return np.ones((N, K), dtype=np.float32)
    return O.make_ternary(N, K, seed=seed, zero_free=(kind == "zero_free"))


def main() -> None:
    ref = O.RefLib()
    OUT.mkdir(parents=True, exist_ok=True)
    for i, (name, type_, K, N, T, wk, ak) in enumerate(SMALL_CASES):
        w = gen_weights(wk, N, K, seed=1000 + i)
        x = O.make_acts(T, K, seed=2000 + i, kind=ak)
        if ak == "outlier":
            x[0] *= 1e-3  # one tiny-magnitude token as well
        packed = ref.pack(type_, w)
        vy, q, sc = ref.quantize(x)
        out1 = ref.gemm(type_, packed, vy, K, N, T, nth=1)
        out4 = ref.gemm(type_, packed, vy, K, N, T, nth=4)
        assert (out1.view(np.uint32) == out4.view(np.uint32)).all(), name
        np.savez_compressed(OUT / f"{name}.npz", type=np.int32(type_), K=np.int64(K), N=np.int64(N), T=np.int64(T),
                            w=w.astype(np.int8), packed=packed, x=x, q=q, scales=sc, out=out1)
        print(f"{name}: {O.TYPE_NAMES[type_]} K={K} N={N} T={T} out[0,:3]={out1[0, :3]}")

    kats = {}
    for name, type_, K, N, T in KAT_CASES:
        w = O.make_ternary(N, K, seed=1234)
        x = O.make_acts(T, K, seed=4321, kind="normal")
        packed = ref.pack(type_, w)
        out = ref.mul_mat(type_, packed, x, N, nth=ref.max_threads())
        kats[name] = {
            "type": type_, "K": K, "N": N, "T": T, "w_seed": 1234, "x_seed": 4321, "x_kind": "normal",
            "sha256_packed": hashlib.sha256(packed.tobytes()).hexdigest(),
            "sha256_x": hashlib.sha256(x.tobytes()).hexdigest(),
            "sha256_out": hashlib.sha256(out.tobytes()).hexdigest(),
            "out_sum_f64": float(out.astype(np.float64).sum()),
        }
        print(name, kats[name]["sha256_out"][:16])
    (OUT / "kat_full.json").write_text(json.dumps(kats, indent=1) + "\n")


if __name__ == "__main__":
    main()
