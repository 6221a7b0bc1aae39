"""oracle/oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes front-ends for the two CPU checkers of the Vec-LUT MUL_MAT path:

* ``Oracle``  -- our plain-C restatement (oracle/vlut_oracle.c -> oracle/_build/libvlut_oracle.so)
* ``RefLib``  -- the reference's OWN kernel sources compiled in place (oracle/Makefile ->
                 oracle/_ref/libvlut_ref_<isa>.so); the strongest oracle and the timed CPU baseline.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
The product path (vlut.cpp_b200/) never does, and fails loudly without its CUDA library.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent

# ggml type ids (reference ggml/include/ggml.h:425-434)
I2_V, I2_V_4, I2_V_8, I1_V, I1_V_2 = 40, 42, 43, 45, 46
TYPE_NAMES = {I2_V: "I2_V", I2_V_4: "I2_V_4", I2_V_8: "I2_V_8", I1_V: "I1_V", I1_V_2: "I1_V_2"}
ALL_TYPES = tuple(TYPE_NAMES)


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def build(ref: bool = True) -> None:
    """Compile the checkers (make -C oracle). The reference part is skipped when /root/reference is absent."""
    subprocess.run(["make", "-C", str(HERE), "oracle"] + (["ref"] if ref else []), check=True,
                   stdout=subprocess.DEVNULL)


def row_size(type_: int, K: int) -> int:
    if type_ in (I1_V, I1_V_2):
        return K // 20 * 4 + K % 20 // 4
    return K // 4


class Oracle:
    """Plain-C restatement (the 'port')."""

    def __init__(self):
        p = HERE / "_build" / "libvlut_oracle.so"
        if not p.exists():
            build(ref=False)
        self.lib = C.CDLL(str(p))
        L = self.lib
        L.vo_row_size.restype = C.c_int64
        L.vo_row_size.argtypes = [C.c_int, C.c_int64]
        L.vo_pack.restype = C.c_int64
        L.vo_pack.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.vo_unpack.restype = C.c_int
        L.vo_unpack.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.vo_quantize_row.restype = None
        L.vo_quantize_row.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]
        L.vo_groups.restype = C.c_int
        L.vo_groups.argtypes = [C.c_int, C.c_int64, C.c_void_p]
        L.vo_mul_mat.restype = C.c_int
        L.vo_mul_mat.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]

    def pack(self, type_: int, w: np.ndarray) -> np.ndarray:
        """w: float32 [N][K] ternary-valued -> packed uint8, whole-matrix GGUF layout."""
        w = np.ascontiguousarray(w, dtype=np.float32)
        N, K = w.shape
        out = np.zeros(N * row_size(type_, K), dtype=np.uint8)
        n = self.lib.vo_pack(type_, _ptr(w), _ptr(out), N, K)
        assert n == out.size
        return out

    def unpack(self, type_: int, packed: np.ndarray, N: int, K: int) -> np.ndarray:
        out = np.zeros((N, K), dtype=np.int8)
        assert self.lib.vo_unpack(type_, _ptr(np.ascontiguousarray(packed)), _ptr(out), N, K) == 0
        return out

    def quantize(self, x: np.ndarray):
        """x: float32 [T][K] -> (int8 q [T][K], float32 scales [T])."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        T, K = x.shape
        q = np.zeros((T, K), dtype=np.int8)
        s = np.zeros(T, dtype=np.float32)
        for t in range(T):
            self.lib.vo_quantize_row(C.c_void_p(x.ctypes.data + t * K * 4), K, C.c_void_p(q.ctypes.data + t * K), 1,
                                     C.c_void_p(s.ctypes.data + 4 * t))
        return q, s

    def groups(self, type_: int, K: int) -> list[int]:
        b = np.zeros(K // 512 + 4, dtype=np.int64)
        ng = self.lib.vo_groups(type_, K, _ptr(b))
        return [int(v) for v in b[: ng + 1]]

    def mul_mat(self, type_: int, packed: np.ndarray, x: np.ndarray, N: int, nth: int = 0, want_acc: bool = False):
        """The full op. Returns out [T][N] float32 (and q, scales, acc32 when want_acc)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        T, K = x.shape
        out = np.zeros((T, N), dtype=np.float32)
        q = np.zeros((T, K), dtype=np.int8)
        s = np.zeros(T, dtype=np.float32)
        acc = np.zeros((T, N), dtype=np.int32)
        nth = nth or os.cpu_count() or 1
        rc = self.lib.vo_mul_mat(type_, K, N, T, _ptr(np.ascontiguousarray(packed)), _ptr(x), K, _ptr(out), _ptr(q),
                                 _ptr(s), _ptr(acc), nth)
        assert rc == 0
        return (out, q, s, acc) if want_acc else out


def _cpu_flags() -> set[str]:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def ref_isa_candidates() -> list[str]:
    f = _cpu_flags()
    v3 = {"avx2", "fma", "bmi2", "movbe", "f16c"} <= f
    v4 = v3 and {"avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl"} <= f
    out = []
    if v4:
        out += ["v4", "v4i"]
    if v3:
        out += ["v3"]
    return out + ["v2"]


class RefLib:
    """The reference's own compiled kernels (kind = 'reference')."""

    def __init__(self, isa: str | None = None, path: Path | None = None, uneven_slices_ok: bool = False):
        """``path``: load this build instead (tests/test_upstream_fixes.py: the kernels with docs/upstream_fixes.patch
        applied, which also lifts the equal-slices restriction of the I1_V_2 tail, ``uneven_slices_ok``)."""
        self.uneven_slices_ok = uneven_slices_ok
        d = HERE / "_ref"
        isa = isa or os.environ.get("VLUT_REF_ISA")
        cands = [isa] if isa else ref_isa_candidates()
        if path is None:
            path = next((d / f"libvlut_ref_{i}.so" for i in cands if (d / f"libvlut_ref_{i}.so").exists()), None)
        if path is None and Path("/root/reference/ggml/src/ggml-quants-vlut.c").exists():
            build(ref=True)
            path = next((d / f"libvlut_ref_{i}.so" for i in cands if (d / f"libvlut_ref_{i}.so").exists()), None)
        if path is None:
            raise FileNotFoundError("oracle/_ref/libvlut_ref_*.so not built (run `make -C oracle ref` where "
                                    "/root/reference exists)")
        self.path = Path(path)
        self.isa = self.path.stem.rsplit("_", 1)[1]
        self.lib = C.CDLL(str(path))
        L = self.lib
        L.vlut_ref_row_size.restype = C.c_size_t
        L.vlut_ref_row_size.argtypes = [C.c_int, C.c_int64]
        L.vlut_ref_pack.restype = C.c_long
        L.vlut_ref_pack.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.vlut_ref_act_bytes.restype = C.c_size_t
        L.vlut_ref_act_bytes.argtypes = [C.c_int64, C.c_int64]
        L.vlut_ref_quantize_act.restype = None
        L.vlut_ref_quantize_act.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int]
        L.vlut_ref_gemm.restype = C.c_int
        L.vlut_ref_gemm.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_int]
        L.vlut_ref_mul_mat.restype = C.c_int
        L.vlut_ref_mul_mat.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                                       C.c_void_p, C.c_int]
        L.vlut_ref_max_threads.restype = C.c_int

    def max_threads(self) -> int:
        return int(self.lib.vlut_ref_max_threads())

    def pack(self, type_: int, w: np.ndarray) -> np.ndarray:
        w = np.ascontiguousarray(w, dtype=np.float32)
        N, K = w.shape
        out = np.zeros(N * row_size(type_, K), dtype=np.uint8)
        n = self.lib.vlut_ref_pack(type_, _ptr(w), _ptr(out), N, K)
        assert n == out.size, (n, out.size)
        return out

    def quantize(self, x: np.ndarray, nth: int = 1):
        """-> (raw tile-layout buffer, int8 q [T][K] row-major view-copy, scales [T])."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        T, K = x.shape
        vy = np.zeros(self.lib.vlut_ref_act_bytes(K, T), dtype=np.uint8)
        self.lib.vlut_ref_quantize_act(_ptr(x), K, K, T, _ptr(vy), nth)
        Tp = (T + 31) // 32 * 32
        tiles = vy[: Tp * K].view(np.int8).reshape(Tp // 32, K, 32)  # [tile][k][t%32]
        q = np.ascontiguousarray(tiles.transpose(0, 2, 1).reshape(Tp, K)[:T])
        sc = vy[Tp * K: Tp * K + 4 * T].view(np.float32).copy()
        return vy, q, sc

    def safe_nth(self, type_: int, K: int, N: int, nth: int) -> int:
        nth = max(1, nth)
        # the stock I1_V_2 tail addresses its weights with `- ith * nc`: only right when every thread has the same nc
        if type_ == I1_V_2 and K % 20 != 0 and not self.uneven_slices_ok:
            while N % nth:
                nth -= 1
        return nth

    def gemm(self, type_: int, packed: np.ndarray, vy: np.ndarray, K: int, N: int, T: int, nth: int = 1) -> np.ndarray:
        out = np.zeros((T, N), dtype=np.float32)
        rc = self.lib.vlut_ref_gemm(type_, K, N, T, _ptr(np.ascontiguousarray(packed)), _ptr(vy), _ptr(out),
                                    self.safe_nth(type_, K, N, nth))
        assert rc == 0, rc
        return out

    def mul_mat(self, type_: int, packed: np.ndarray, x: np.ndarray, N: int, nth: int = 1) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        T, K = x.shape
        out = np.zeros((T, N), dtype=np.float32)
        rc = self.lib.vlut_ref_mul_mat(type_, K, N, T, _ptr(np.ascontiguousarray(packed)), _ptr(x), K, _ptr(out),
                                       self.safe_nth(type_, K, N, nth))
        assert rc == 0, rc
        return out


# --------------------------------------------------------------------------------------------------------------
# synthetic inputs: deterministic, numpy PCG64 so they can be regenerated anywhere
# --------------------------------------------------------------------------------------------------------------
def make_ternary(N: int, K: int, seed: int = 1234, zero_free: bool = False) -> np.ndarray:
    rng = np.random.Generator(np.random.PCG64(seed))
    if zero_free:  # what tests/test-vlut-gemm.cpp's uniform(-1,1) init yields: never 0
        return (rng.integers(0, 2, size=(N, K), dtype=np.int8) * 2 - 1).astype(np.float32)
    return (rng.integers(0, 3, size=(N, K), dtype=np.int8) - 1).astype(np.float32)


def make_acts(T: int, K: int, seed: int = 4321, kind: str = "normal") -> np.ndarray:
    rng = np.random.Generator(np.random.PCG64(seed))
    if kind == "normal":
        x = rng.standard_normal((T, K), dtype=np.float32)
    elif kind == "uniform":
        x = rng.uniform(-1.0, 1.0, size=(T, K)).astype(np.float32)
    elif kind == "outlier":  # a few large-magnitude channels, like real LLM activations
        x = rng.standard_normal((T, K), dtype=np.float32)
        cols = rng.integers(0, K, size=max(1, K // 256))
        x[:, cols] *= 40.0
    elif kind == "negative":  # all-negative rows: signed-max rule saturates everything to -128
        x = -np.abs(rng.standard_normal((T, K), dtype=np.float32)) - 0.5
    elif kind == "ones":  # adversarial int16 wrap with all-(+1) weights
        x = np.ones((T, K), dtype=np.float32)
    elif kind == "zeros":
        x = np.zeros((T, K), dtype=np.float32)
    else:
        raise ValueError(kind)
    return x
