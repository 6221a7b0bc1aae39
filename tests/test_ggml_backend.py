"""The ggml backend plugin (libggml-vlut-b200.so) exercised through the reference's own ggml API.

oracle/_ref/{libggml-base,libggml,libggml-cpu}.so are the reference's UNMODIFIED ggml sources compiled in place by
oracle/Makefile; tests/harness/ggml_harness.cpp (-> oracle/_ref/vlut_ggml_harness) is a small driver that loads the
plugin with ggml_backend_load() and runs GGML_OP_MUL_MAT graphs the way the reference's tools do.
"""
import ctypes
import os
import subprocess
from pathlib import Path

import numpy as np
import pytest

from oracle import oracle as O

ROOT = Path(__file__).resolve().parent.parent
REFDIR = ROOT / "oracle" / "_ref"
HARNESS = REFDIR / "vlut_ggml_harness"
PLUGIN = ROOT / "vlut.cpp_b200" / "csrc" / "build" / "libggml-vlut-b200.so"
CPU_BACKEND = REFDIR / "libggml-cpu.so"


def need_artifacts():
    missing = [p for p in (HARNESS, PLUGIN, CPU_BACKEND) if not p.exists()]
    if missing:
        pytest.skip(f"not built (needs /root/reference at build time): {missing}")


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def run_harness(tmp_path, mode, type_, K, N, T, packed, x, env=None):
    (tmp_path / "w.bin").write_bytes(packed.tobytes())
    (tmp_path / "x.bin").write_bytes(np.ascontiguousarray(x, dtype=np.float32).tobytes())
    cmd = [str(HARNESS), str(PLUGIN), mode, str(type_), str(K), str(N), str(T), str(tmp_path / "w.bin"), str(tmp_path / "x.bin"),
           str(tmp_path / "out.bin")]
    if mode in ("sched", "block"):
        cmd.append(str(CPU_BACKEND))
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120,
                       env={**os.environ, **(env or {})})
    return r


def test_plugin_exports_ggml_entry_points():
    need_artifacts()
    # resolve against libggml-base like the host process would
    ctypes.CDLL(str(REFDIR / "libggml-base.so"), mode=ctypes.RTLD_GLOBAL)
    lib = ctypes.CDLL(str(PLUGIN))
    assert hasattr(lib, "ggml_backend_init") and hasattr(lib, "ggml_backend_score")
    lib.ggml_backend_score.restype = ctypes.c_int
    import torch
    score = lib.ggml_backend_score()
    assert (score > 0) == torch.cuda.is_available()


def test_plugin_is_refused_without_gpu(tmp_path, orc):
    """ggml_backend_load() asks ggml_backend_score(): 0 on a box without an sm_100 GPU -> not loaded, no CPU fallback."""
    need_artifacts()
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    w = O.make_ternary(16, 512, seed=1)
    r = run_harness(tmp_path, "direct", O.I2_V_4, 512, 16, 2, orc.pack(O.I2_V_4, w), O.make_acts(2, 512))
    assert r.returncode == 3, r.stdout
    assert "failed" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("type_,K,N,T", [(O.I2_V_4, 4096, 512, 8), (O.I2_V_4, 2048, 300, 70), (O.I1_V_2, 3200, 384, 33),
                                         (O.I2_V, 1024, 128, 1), (O.I2_V_8, 2048, 256, 130), (O.I1_V, 1300, 64, 17),
                                         (O.I1_V_2, 4096, 128, 5)])
def test_mul_mat_through_ggml_direct(tmp_path, orc, type_, K, N, T):
    need_artifacts()
    w = O.make_ternary(N, K, seed=K + T)
    x = O.make_acts(T, K, seed=N, kind="outlier")
    packed = orc.pack(type_, w)
    r = run_harness(tmp_path, "direct", type_, K, N, T, packed, x)
    assert r.returncode == 0, r.stdout
    assert "VLUT_B200" in r.stdout
    out = np.fromfile(tmp_path / "out.bin", dtype=np.float32).reshape(T, N)
    assert np.array_equal(bits(out), bits(orc.mul_mat(type_, packed, x, N)))


@pytest.mark.gpu
@pytest.mark.parametrize("type_,K,N,T", [(O.I2_V_4, 4096, 512, 16), (O.I1_V_2, 3200, 256, 40), (O.I2_V_4, 2048, 1024, 512)])
def test_mul_mat_through_ggml_sched(tmp_path, orc, type_, K, N, T):
    """Weight in the plugin's buffer, everything else on the CPU backend: ggml_backend_sched must route the MUL_MAT to
    the plugin (supports_op / supports_buft) and copy activations in and results out around the split."""
    need_artifacts()
    w = O.make_ternary(N, K, seed=K + T)
    x = O.make_acts(T, K, seed=N)
    packed = orc.pack(type_, w)
    r = run_harness(tmp_path, "sched", type_, K, N, T, packed, x)
    assert r.returncode == 0, r.stdout
    assert "mul_mat ran on 'VLUT_B200_0'" in r.stdout, r.stdout
    out = np.fromfile(tmp_path / "out.bin", dtype=np.float32).reshape(T, N)
    expect = orc.mul_mat(type_, packed, x, N) + np.float32(1.0)
    assert np.array_equal(bits(out), bits(expect))


@pytest.mark.gpu
@pytest.mark.parametrize("fuse", [True, False], ids=["fused", "VLUT_B200_NO_FUSE"])
@pytest.mark.parametrize("type_,K,N,T", [(O.I2_V_4, 4096, 512, 16), (O.I1_V_2, 3200, 384, 40), (O.I2_V_4, 2048, 1024, 200)])
def test_scale_fusion_and_shared_activations(tmp_path, orc, type_, K, N, T, fuse):
    """SURVEY 8f N1 + N3: the `.scale` MUL that follows a ternary MUL_MAT (src/models/bitnet.cpp:33-34) is folded into
    the GEMM epilogue when the product has one consumer, runs as a stand-alone kernel otherwise, and four MUL_MATs on
    one input quantise it once.  Bits must equal MUL_MAT (oracle) followed by an fp32 multiply, fused or not."""
    need_artifacts()
    w = O.make_ternary(N, K, seed=K + T)
    x = O.make_acts(T, K, seed=N, kind="outlier")
    packed = orc.pack(type_, w)
    r = run_harness(tmp_path, "block", type_, K, N, T, packed, x, env=None if fuse else {"VLUT_B200_NO_FUSE": "1"})
    assert r.returncode == 0, r.stdout
    assert "scaled product ran on 'VLUT_B200_0'" in r.stdout, r.stdout
    out = np.fromfile(tmp_path / "out.bin", dtype=np.float32).reshape(5, T, N)
    base = orc.mul_mat(type_, packed, x, N)
    expect = [base * np.float32(0.37), base * np.float32(-1.9), base, base * np.float32(0.37), base + np.float32(1.0)]
    for i, e in enumerate(expect):
        assert np.array_equal(bits(out[i]), bits(e)), f"output {i} differs"
    # fused: the four MUL_MATs on one input (two of them with their scale folded in) run as ONE batched group
    want = ("gemm=4 quantize=1 fused_scale=2 scale_rows=1 batched=4" if fuse
            else "gemm=4 quantize=1 fused_scale=0 scale_rows=3 batched=2")
    assert want in r.stdout, r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["direct", "sched", "block"])
@pytest.mark.parametrize("gpus", [2, 3])
def test_one_logical_device_owning_several_gpus(tmp_path, orc, mode, gpus):
    """VLUT_B200_GPUS=G: ONE ggml device whose set_tensor row-shards every vlut weight over G GPUs and whose graph_compute
    runs the MUL_MAT on all of them with the gather fused into the GEMM epilogues (SURVEY.md section 8e "Exposure").  On a box
    with fewer GPUs the members share GPU 0 (VLUT_B200_GROUP_SAME_DEVICE): same code path."""
    need_artifacts()
    import torch
    env = {"VLUT_B200_GPUS": str(gpus)}
    if torch.cuda.device_count() < gpus:
        env["VLUT_B200_GROUP_SAME_DEVICE"] = "1"
    type_, K, N, T = O.I2_V_4, 2048, 1000, 70
    w = O.make_ternary(N, K, seed=gpus)
    x = O.make_acts(T, K, seed=N, kind="outlier")
    packed = orc.pack(type_, w)
    r = run_harness(tmp_path, mode, type_, K, N, T, packed, x, env=env)
    assert r.returncode == 0, r.stdout
    assert f"row-sharded over {gpus} GPUs" in r.stdout, r.stdout
    base = orc.mul_mat(type_, packed, x, N)
    if mode == "direct":
        out = np.fromfile(tmp_path / "out.bin", dtype=np.float32).reshape(T, N)
        assert np.array_equal(bits(out), bits(base))
    elif mode == "sched":
        out = np.fromfile(tmp_path / "out.bin", dtype=np.float32).reshape(T, N)
        assert np.array_equal(bits(out), bits(base + np.float32(1.0)))
    else:
        out = np.fromfile(tmp_path / "out.bin", dtype=np.float32).reshape(5, T, N)
        expect = [base * np.float32(0.37), base * np.float32(-1.9), base, base * np.float32(0.37), base + np.float32(1.0)]
        for i, e in enumerate(expect):
            assert np.array_equal(bits(out[i]), bits(e)), f"output {i} differs"
        assert "sharded=4" in r.stdout, r.stdout
