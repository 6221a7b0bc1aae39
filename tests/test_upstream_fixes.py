"""docs/upstream_fixes.patch: every defect it fixes is reproduced on the STOCK reference build and shown gone on the patched
one.  CPU only; needs the artefacts scripts/verify_upstream_fixes.sh builds from /root/reference (oracle/_ref/upfix/, built by
__graft_entry__.build() in the build container) -- skipped where they do not exist."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

from oracle import oracle as O

ROOT = Path(__file__).resolve().parent.parent
UPFIX = ROOT / "oracle" / "_ref" / "upfix"


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def need():
    miss = [p for p in (UPFIX / "stock" / "vlut_cpu_mul_mat", UPFIX / "patched" / "vlut_cpu_mul_mat",
                        UPFIX / "libvlut_ref_patched.so") if not p.exists()]
    if miss:
        pytest.skip(f"not built (scripts/verify_upstream_fixes.sh needs /root/reference): {miss[0].name}")


@pytest.fixture(scope="module")
def orc():
    return O.Oracle()


def cpu_mul_mat(build, tmp_path, type_, K, N, T, packed, x, threads=4):
    (tmp_path / "w.bin").write_bytes(packed.tobytes())
    (tmp_path / "x.bin").write_bytes(np.ascontiguousarray(x, dtype=np.float32).tobytes())
    r = subprocess.run([str(UPFIX / build / "vlut_cpu_mul_mat"), str(type_), str(K), str(N), str(T), str(tmp_path / "w.bin"),
                        str(tmp_path / "x.bin"), str(tmp_path / "out.bin"), str(threads)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0, r.stdout
    out = np.fromfile(tmp_path / "out.bin", dtype=np.float32).reshape(T, N)
    sizes = dict(kv.split("=") for kv in r.stdout.split() if "=" in kv)
    return out, int(sizes["work_size"]), int(sizes["vlut_path_writes"])


@pytest.mark.parametrize("T", [32, 33, 40, 96])
def test_fix1_activation_tile_base(tmp_path, orc, T):
    """ggml-cpu.c:1404: beyond the first 32-token tile the stock wrapper quantises into the wrong place."""
    need()
    type_, K, N = O.I2_V_4, 1024, 64
    w = O.make_ternary(N, K, seed=11)
    x = O.make_acts(T, K, seed=12)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    stock, _, _ = cpu_mul_mat("stock", tmp_path, type_, K, N, T, packed, x)
    fixed, _, _ = cpu_mul_mat("patched", tmp_path, type_, K, N, T, packed, x)
    assert np.array_equal(bits(fixed), bits(ref))
    if T <= 32:
        assert np.array_equal(bits(stock), bits(ref))
    else:
        assert not np.array_equal(bits(stock), bits(ref)), "stock wrapper is right for T > 32: drop fix 1 from the patch"


@pytest.mark.parametrize("type_", [O.I1_V, O.I1_V_2])
@pytest.mark.parametrize("K", [4096, 1304, 656])
def test_fix2_i1_tail_table(orc, type_, K):
    """ggml-cpu-quants-vlut.c:1570 / :1601 / :1678 / :1713: K % 20 != 0 tails look their I2 blocks up in a table built
    on a stale entry.  Patched kernels == intended integer GEMM (the oracle); stock kernels differ."""
    need()
    assert K % 20 != 0
    N, T = 64, 40
    w = O.make_ternary(N, K, seed=K)
    x = O.make_acts(T, K, seed=K + 1)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    stock = O.RefLib().mul_mat(type_, packed, x, N, nth=2)
    fixed = O.RefLib(path=UPFIX / "libvlut_ref_patched.so", uneven_slices_ok=True).mul_mat(type_, packed, x, N, nth=2)
    assert np.array_equal(bits(fixed), bits(ref))
    assert not np.array_equal(bits(stock), bits(ref)), "stock tail is right: drop fix 2 from the patch"


def test_fix2_i1v2_tail_with_uneven_thread_slices(orc):
    """... and the I1_V_2 tail's `- ith * nc` weight offset: N = 66 rows on 4 threads (slices of 16, 17, 16, 17)."""
    need()
    type_, K, N, T = O.I1_V_2, 1304, 66, 8
    w = O.make_ternary(N, K, seed=3)
    x = O.make_acts(T, K, seed=4)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    fixed = O.RefLib(path=UPFIX / "libvlut_ref_patched.so", uneven_slices_ok=True).mul_mat(type_, packed, x, N, nth=4)
    assert np.array_equal(bits(fixed), bits(ref))


@pytest.mark.parametrize("K,T,stock_fits", [(2048, 1, True), (14336, 1, True), (16384, 1, False), (20480, 33, False)])
def test_fix3_work_buffer_size(tmp_path, orc, K, T, stock_fits):
    """ggml-cpu.c:2943-2949 + :3074-3077: the plan reserves ne10 * ne11 + 4 bytes plus a fixed slack of 32 * 15000 + 2048 * 4
    bytes; the path writes ceil32(ne11) * ne10 + 4 * ne11.  The slack covers K <= ~15000 (every released model); beyond it
    a small-T MUL_MAT overruns the work buffer.  The patched plan reserves what the path writes."""
    need()
    type_, N = O.I2_V_4, 32
    w = O.make_ternary(N, K, seed=1)
    x = O.make_acts(T, K, seed=2)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    _, ws_stock, need_bytes = cpu_mul_mat("stock", tmp_path, type_, K, N, T, packed, x, threads=1)
    fixed, ws_fixed, need2 = cpu_mul_mat("patched", tmp_path, type_, K, N, T, packed, x, threads=1)
    assert need_bytes == need2 == (T + 31) // 32 * 32 * K + 4 * T
    assert (ws_stock >= need_bytes) == stock_fits
    assert ws_fixed >= need_bytes
    assert np.array_equal(bits(fixed), bits(ref))
