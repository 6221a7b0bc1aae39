"""CPU suite, part 1: pin the oracle.

* the plain-C restatement (oracle/vlut_oracle.c) against every committed golden fixture -- each fixture is an
  output of the REFERENCE's own compiled kernels (oracle/make_golden.py);
* the restatement against the compiled reference itself (oracle/_ref) on fresh seeded inputs, thread-sliced
  like ggml-cpu.c:1413-1418, including the int16-wrap and the signed-max edge cases;
* the documented deviation of the reference for I1 types with K % 20 != 0.
"""
import glob
import hashlib
import json
from pathlib import Path

import numpy as np
import pytest

from oracle import oracle as O

GOLDEN = Path(__file__).resolve().parent / "golden"
FIXTURES = sorted(glob.glob(str(GOLDEN / "*.npz")))


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def test_fixtures_present():
    assert len(FIXTURES) >= 10
    assert (GOLDEN / "kat_full.json").exists()


@pytest.mark.parametrize("path", FIXTURES, ids=lambda p: Path(p).stem)
def test_oracle_matches_golden(orc, path):
    z = np.load(path)
    type_, K, N, T = int(z["type"]), int(z["K"]), int(z["N"]), int(z["T"])
    # packer (ggml-quants-vlut.c) and its inverse
    packed = orc.pack(type_, z["w"].astype(np.float32))
    assert np.array_equal(packed, z["packed"])
    assert np.array_equal(orc.unpack(type_, z["packed"], N, K), z["w"])
    # activation quantiser (ggml-cpu-quants-vlut.c:952-973)
    q, s = orc.quantize(z["x"])
    assert np.array_equal(q, z["q"])
    assert np.array_equal(bits(s), bits(z["scales"]))
    # the op, bit-exact on fp32
    out = orc.mul_mat(type_, z["packed"], z["x"], N)
    assert out.shape == (T, N)
    assert np.array_equal(bits(out), bits(z["out"]))


CASES = [
    # type, K, N, T
    (O.I2_V, 1024, 40, 33), (O.I2_V, 516, 24, 5),
    (O.I2_V_4, 4096, 64, 8), (O.I2_V_4, 1040, 32, 65), (O.I2_V_4, 512, 8, 1),
    (O.I2_V_8, 2048, 32, 31), (O.I2_V_8, 1056, 16, 34),
    (O.I1_V, 1280, 48, 33), (O.I1_V, 660, 16, 3), (O.I1_V, 3200, 16, 32),
    (O.I1_V_2, 640, 64, 70), (O.I1_V_2, 3220, 24, 9), (O.I1_V_2, 8640, 8, 2),
]


@pytest.mark.parametrize("type_,K,N,T", CASES)
@pytest.mark.parametrize("kind", ["normal", "uniform", "outlier"])
def test_oracle_matches_reference_kernels(orc, reflib, type_, K, N, T, kind):
    w = O.make_ternary(N, K, seed=K * 7 + N)
    x = O.make_acts(T, K, seed=T * 13 + K, kind=kind)
    packed = reflib.pack(type_, w)
    assert np.array_equal(packed, orc.pack(type_, w))
    _, q_ref, s_ref = reflib.quantize(x)
    out, q, s, acc = orc.mul_mat(type_, packed, x, N, want_acc=True)
    assert np.array_equal(q, q_ref) and np.array_equal(bits(s), bits(s_ref))
    for nth in (1, 4):
        ref = reflib.mul_mat(type_, packed, x, N, nth=nth)
        assert np.array_equal(bits(out), bits(ref)), f"nth={nth}"
    # the int32 accumulator is consistent with the fp32 result
    assert np.array_equal(bits(out), bits(acc.astype(np.float32) * s[:, None]))


@pytest.mark.parametrize("type_,K,expect", [(O.I2_V_4, 4096, -32.25197), (O.I1_V_2, 3200, 619.8425)])
def test_int16_wrap_is_part_of_the_spec(orc, reflib, type_, K, expect):
    # all-(+1) weights x all-127 activations: every K-group overflows int16
    w = np.ones((8, K), dtype=np.float32)
    x = O.make_acts(2, K, kind="ones")
    packed = reflib.pack(type_, w)
    ref = reflib.mul_mat(type_, packed, x, 8)
    out = orc.mul_mat(type_, packed, x, 8)
    assert np.array_equal(bits(ref), bits(out))
    assert abs(float(out[0, 0]) - expect) < 1e-3


def test_signed_max_rule(orc, reflib):
    # all-negative token: max(1e-5, max x) = 1e-5 -> everything saturates to -128, scale = 1e-5/127
    x = O.make_acts(3, 512, kind="negative")
    q, s = orc.quantize(x)
    assert (q == -128).all()
    assert np.array_equal(bits(s), bits(np.full(3, np.float32(1e-5 / 127))))
    _, q_ref, s_ref = reflib.quantize(x)
    assert np.array_equal(q, q_ref) and np.array_equal(bits(s), bits(s_ref))


def test_group_boundaries(orc):
    assert orc.groups(O.I2_V_4, 4096) == [0, 512, 1024, 1536, 2048, 2560, 3072, 3584, 4096]
    assert orc.groups(O.I2_V, 1040) == [0, 512, 1024, 1040]
    assert orc.groups(O.I1_V_2, 3200) == [0, 640, 1280, 1920, 2560, 3200]
    assert orc.groups(O.I1_V, 4096) == [0, 640, 1280, 1920, 2560, 3200, 3840, 4096]
    assert orc.groups(O.I1_V, 640) == [0, 640]


@pytest.mark.parametrize("type_", [O.I1_V, O.I1_V_2])
def test_reference_i1_tail_deviation_is_documented(orc, reflib, type_):
    """For K % 20 != 0 the reference's tail table is built on a stale zero entry
    (ggml-cpu-quants-vlut.c:1570-1572 clears entry 0, gemm_make_table_i2v builds from entry 40): its result
    differs from its own intended math.  The spec (and the oracle) is the intended math; this test keeps the
    deviation visible so that it is re-examined if the reference is ever fixed."""
    K, N, T = 4096, 64, 8
    w = O.make_ternary(N, K, seed=5)
    x = O.make_acts(T, K, seed=6)
    packed = reflib.pack(type_, w)
    ref = reflib.mul_mat(type_, packed, x, N)
    out = orc.mul_mat(type_, packed, x, N)
    frac = float(np.mean(bits(ref) != bits(out)))
    assert frac > 0.2, "reference now agrees with the intended math for K%20!=0: update SURVEY/DESIGN"
    # ... and the intended math is what a plain integer GEMM on the decoded weights gives
    q, s = orc.quantize(x)
    acc = q.astype(np.int32) @ w.astype(np.int32).T
    assert np.array_equal(bits(out), bits(acc.astype(np.float32) * s[:, None]))


def test_kat_inputs_reproducible():
    """The full-size known answers (tests/golden/kat_full.json) are keyed on seeded numpy inputs: make sure this
    numpy build regenerates exactly those inputs (otherwise the GPU KAT test would be meaningless)."""
    kats = json.loads((GOLDEN / "kat_full.json").read_text())
    k = kats["bitnet_i1v2_3200x8640x65"]
    x = O.make_acts(k["T"], k["K"], seed=k["x_seed"], kind=k["x_kind"])
    assert hashlib.sha256(x.tobytes()).hexdigest() == k["sha256_x"]


def test_kat_smallest_via_oracle(orc):
    kats = json.loads((GOLDEN / "kat_full.json").read_text())
    k = kats["bitnet_i1v2_3200x8640x65"]
    w = O.make_ternary(k["N"], k["K"], seed=k["w_seed"])
    x = O.make_acts(k["T"], k["K"], seed=k["x_seed"], kind=k["x_kind"])
    packed = orc.pack(k["type"], w)
    assert hashlib.sha256(packed.tobytes()).hexdigest() == k["sha256_packed"]
    out = orc.mul_mat(k["type"], packed, x, k["N"])
    assert hashlib.sha256(out.tobytes()).hexdigest() == k["sha256_out"]
