"""Host-side mirror of the reference's operator interface for the Vec-LUT MUL_MAT path, over the C ABI.

Names follow the reference (ggml): weight *types* I2_V / I2_V_4 / I2_V_8 / I1_V / I1_V_2 with ggml's enum values
(ggml/include/ggml.h:425-434), ``nbytes``/``row_size`` like ggml_nbytes / ggml_row_size (ggml/src/ggml.c:1286-1335),
``mul_mat(weight, x)`` like ggml_mul_mat(a, b) on a vlut-typed ``a`` (ggml/src/ggml-cpu/ggml-cpu.c:1372-1421):
x is f32 [T][K] (ggml ne = {K, T}), the result f32 [T][N] (ggml ne = {N, T}).

torch is used for device memory and streams only; all arithmetic runs in libvlut_b200.so.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from ._lib import check

# enum ggml_type values
I2_V, I2_V_4, I2_V_8, I1_V, I1_V_2 = 40, 42, 43, 45, 46
TYPE_NAMES = {I2_V: "I2_V", I2_V_4: "I2_V_4", I2_V_8: "I2_V_8", I1_V: "I1_V", I1_V_2: "I1_V_2"}
TYPE_BY_NAME = {v: k for k, v in TYPE_NAMES.items()}

ALGO_AUTO, ALGO_SIMPLE, ALGO_LUT, ALGO_TCP, ALGO_TCS, ALGO_GEMV = 0, 1, 4, 7, 8, 9
ALGO_NAMES = {ALGO_AUTO: "auto", ALGO_SIMPLE: "simple", ALGO_LUT: "lut", ALGO_TCP: "tcp", ALGO_TCS: "tcs", ALGO_GEMV: "gemv"}


def reload_env() -> None:
    """Make the library re-read its VLUT_* environment knobs (they are resolved once, never on the launch path)."""
    check(_lib.lib().vlut_reload_env(), "vlut_reload_env")


def op_supported(type_: int, K: int, N: int, T: int) -> bool:
    return bool(_lib.lib().vlut_op_supported(type_, K, N, T))


def nbytes(type_: int, K: int, N: int) -> int:
    """ggml_nbytes of a vlut weight tensor with ne = {K, N}."""
    return int(_lib.lib().vlut_gguf_nbytes(type_, K, N))


def repacked_nbytes(type_: int, K: int, N: int) -> int:
    return int(_lib.lib().vlut_repacked_nbytes(type_, K, N))


def pick_algo(type_: int, K: int, N: int, T: int) -> int:
    return int(_lib.lib().vlut_pick_algo(type_, K, N, T))


def device_count() -> int:
    n = C.c_int(0)
    check(_lib.lib().vlut_device_count(C.byref(n)), "vlut_device_count")
    return n.value


def shard_rows(type_: int, K: int, N: int, row0: int, rows: int, gguf: np.ndarray) -> np.ndarray:
    """Rows [row0, row0+rows) of a GGUF-packed matrix, still GGUF-packed (host-side, for row-sharded TP)."""
    gguf = np.ascontiguousarray(gguf, dtype=np.uint8)
    assert gguf.size == nbytes(type_, K, N), (gguf.size, nbytes(type_, K, N))
    out = np.empty(nbytes(type_, K, rows), dtype=np.uint8)
    check(_lib.lib().vlut_shard_rows_host(type_, K, N, row0, rows, gguf.ctypes.data, out.ctypes.data),
          "vlut_shard_rows_host")
    return out


@dataclass
class Weight:
    """A vlut weight resident on the device in the tiled T2 layout (the backend buffer's view of a tensor)."""
    type: int
    K: int
    N: int
    t2: torch.Tensor  # uint8, cuda

    @property
    def gguf_nbytes(self) -> int:
        return nbytes(self.type, self.K, self.N)


class Context:
    """One device + one stream + workspace (vlut_ctx). By default it rides torch's current stream so that
    torch.cuda.Event timing and torch tensors interoperate without extra synchronisation."""

    def __init__(self, device: int | torch.device | None = None, stream: torch.cuda.Stream | None = None):
        if not torch.cuda.is_available():
            raise RuntimeError("vlut_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("vlut_b200 contexts live on CUDA devices only")
        self.device = torch.device("cuda", dev.index if dev.index is not None else torch.cuda.current_device())
        with torch.cuda.device(self.device):
            self.stream = stream or torch.cuda.current_stream()
        h = C.c_void_p()
        # torch's default stream has handle 0; the C ABI reserves NULL for "create a private stream", so the legacy
        # default stream is named by CUDA's own alias cudaStreamLegacy (0x1)
        handle = self.stream.cuda_stream or 1
        check(_lib.lib().vlut_ctx_create(self.device.index, C.c_void_p(handle), C.byref(h)), "vlut_ctx_create")
        self._h = h

    OPT_INPUTS_VISIBLE = 1

    def set_option(self, option: int, value: int) -> None:
        """vlut_ctx_set_option.  OPT_INPUTS_VISIBLE: the caller promises that activations are already visible on the
        device when passed in (not written by a foreign kernel launched just before on this stream)."""
        check(_lib.lib().vlut_ctx_set_option(self._h, option, value), "vlut_ctx_set_option")

    def close(self) -> None:
        if getattr(self, "_h", None):
            _lib.lib().vlut_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- weights ---------------------------------------------------------------------------------------
    def upload(self, type_: int, K: int, N: int, gguf) -> Weight:
        """GGUF-packed bytes (numpy uint8 on the host, or a torch uint8 tensor already on this device) -> Weight."""
        L = _lib.lib()
        nb = nbytes(type_, K, N)
        if nb == 0:
            raise ValueError(f"invalid vlut tensor shape type={type_} K={K} N={N}")
        with torch.cuda.device(self.device):
            t2 = torch.empty(repacked_nbytes(type_, K, N), dtype=torch.uint8, device=self.device)
            if isinstance(gguf, torch.Tensor) and gguf.is_cuda:
                assert gguf.dtype == torch.uint8 and gguf.numel() == nb and gguf.is_contiguous()
                check(L.vlut_repack_weights(self._h, type_, K, N, gguf.data_ptr(), t2.data_ptr()), "vlut_repack_weights")
            else:
                g = np.ascontiguousarray(gguf.numpy() if isinstance(gguf, torch.Tensor) else gguf, dtype=np.uint8)
                assert g.size == nb, (g.size, nb)
                check(L.vlut_upload_weights(self._h, type_, K, N, g.ctypes.data, t2.data_ptr()), "vlut_upload_weights")
        return Weight(type_, K, N, t2)

    def unpack(self, w: Weight) -> np.ndarray:
        """T2 -> ternary int8 [N][K] on the host (test hook)."""
        out = np.empty((w.N, w.K), dtype=np.int8)
        check(_lib.lib().vlut_debug_unpack_t2(self._h, w.type, w.K, w.N, w.t2.data_ptr(), out.ctypes.data),
              "vlut_debug_unpack_t2")
        return out

    # ---- the op ----------------------------------------------------------------------------------------
    def quantize_act(self, type_: int, x: torch.Tensor) -> None:
        assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
        T, K = x.shape
        check(_lib.lib().vlut_quantize_act(self._h, type_, x.data_ptr(), x.stride(0), K, T), "vlut_quantize_act")

    def get_act(self, K: int, T: int):
        q = np.empty((T, K), dtype=np.int8)
        s = np.empty(T, dtype=np.float32)
        check(_lib.lib().vlut_debug_get_act(self._h, K, T, q.ctypes.data, s.ctypes.data), "vlut_debug_get_act")
        return q, s

    def gemm(self, w: Weight, T: int, out: torch.Tensor | None = None, w_scale: float = 1.0,
             algo: int = ALGO_AUTO) -> torch.Tensor:
        if out is None:
            out = torch.empty((T, w.N), dtype=torch.float32, device=self.device)
        assert out.is_cuda and out.dtype == torch.float32 and out.shape == (T, w.N) and out.stride(1) == 1
        check(_lib.lib().vlut_gemm(self._h, w.type, w.K, w.N, T, w.t2.data_ptr(), out.data_ptr(), out.stride(0),
                                   w_scale, algo), "vlut_gemm")
        return out

    def mul_mat(self, w: Weight, x: torch.Tensor, out: torch.Tensor | None = None, w_scale: float = 1.0,
                algo: int = ALGO_AUTO) -> torch.Tensor:
        """GGML_OP_MUL_MAT(w, x): x f32 [T][K] on the device -> f32 [T][N]."""
        assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1 and x.shape[1] == w.K
        T = x.shape[0]
        if out is None:
            out = torch.empty((T, w.N), dtype=torch.float32, device=self.device)
        assert out.is_cuda and out.dtype == torch.float32 and out.shape == (T, w.N) and out.stride(1) == 1
        check(_lib.lib().vlut_mul_mat(self._h, w.type, w.K, w.N, T, w.t2.data_ptr(), x.data_ptr(), x.stride(0),
                                      out.data_ptr(), out.stride(0), w_scale, algo), "vlut_mul_mat")
        return out

    def mul_mat_batch(self, ws: list[Weight], x: torch.Tensor, outs: list[torch.Tensor] | None = None,
                      w_scales: list[float] | None = None, algo: int = ALGO_AUTO) -> list[torch.Tensor]:
        """Several weights on ONE activation tensor (Q/K/V, up/gate): one quantisation, one launch at small T."""
        assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
        T, K = x.shape
        n = len(ws)
        assert n >= 1 and all(w.K == K for w in ws)
        if outs is None:
            outs = [torch.empty((T, w.N), dtype=torch.float32, device=self.device) for w in ws]
        types = (C.c_int * n)(*[w.type for w in ws])
        Ns = (C.c_int64 * n)(*[w.N for w in ws])
        t2 = (C.c_void_p * n)(*[w.t2.data_ptr() for w in ws])
        op = (C.c_void_p * n)(*[o.data_ptr() for o in outs])
        ldos = (C.c_int64 * n)(*[o.stride(0) for o in outs])
        sc = (C.c_float * n)(*(w_scales or [1.0] * n))
        check(_lib.lib().vlut_mul_mat_batch(self._h, n, types, K, Ns, T, t2, x.data_ptr(), x.stride(0), op, ldos, sc, algo),
              "vlut_mul_mat_batch")
        return outs

    def mul_mat_scatter(self, w: Weight, x: torch.Tensor, peers: list[int], multicast_ptr: int, ldo: int, col0: int,
                        w_scale: float = 1.0, algo: int = ALGO_AUTO) -> None:
        """Row-sharded op with the all-gather fused into the epilogue: this rank's [T][w.N] slice is stored at column
        ``col0`` of the [T][ldo] result buffer of EVERY rank (``peers`` = base addresses of those buffers, or the
        NVSwitch multicast address of the symmetric allocation)."""
        assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1 and x.shape[1] == w.K
        T = x.shape[0]
        if multicast_ptr:
            ptrs = [multicast_ptr + 4 * col0]
        else:
            ptrs = [p + 4 * col0 for p in peers]
        arr = (C.c_void_p * len(ptrs))(*ptrs)
        check(_lib.lib().vlut_mul_mat_scatter(self._h, w.type, w.K, w.N, T, w.t2.data_ptr(), x.data_ptr(), x.stride(0), arr,
                                              len(ptrs), 1 if multicast_ptr else 0, ldo, w_scale, algo),
              "vlut_mul_mat_scatter")

    def mul_mat_scatter_push(self, w: Weight, x: torch.Tensor, peers: list[int], self_index: int, ldo: int, col0: int,
                             w_scale: float = 1.0) -> None:
        """vlut_mul_mat_scatter_push: local GEMM + a concurrent copy kernel that forwards finished tiles to the peers."""
        assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1 and x.shape[1] == w.K
        T = x.shape[0]
        ptrs = [p + 4 * col0 for p in peers]
        arr = (C.c_void_p * len(ptrs))(*ptrs)
        check(_lib.lib().vlut_mul_mat_scatter_push(self._h, w.type, w.K, w.N, T, w.t2.data_ptr(), x.data_ptr(), x.stride(0),
                                                   arr, len(ptrs), self_index, ldo, w_scale), "vlut_mul_mat_scatter_push")

    def barrier(self, flag_ptrs: list[int], rank: int) -> None:
        """vlut_barrier: publish a fused all-gather across ranks (flag_ptrs[i] = rank i's peer-visible flag array)."""
        arr = (C.c_void_p * len(flag_ptrs))(*flag_ptrs)
        check(_lib.lib().vlut_barrier(self._h, arr, len(flag_ptrs), rank), "vlut_barrier")

    def scale_rows(self, x: torch.Tensor, scale: float, out: torch.Tensor | None = None) -> torch.Tensor:
        """out[t][n] = x[t][n] * scale (one fp32 rounding): GGML_OP_MUL by a one-element tensor, the weight's `.scale`
        (ggml-cpu/binary-ops.cpp:152), for when it is not folded into the GEMM epilogue.  In place when out is x."""
        assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
        T, N = x.shape
        if out is None:
            out = torch.empty((T, N), dtype=torch.float32, device=x.device)
        assert out.is_cuda and out.dtype == torch.float32 and out.shape == (T, N) and out.stride(1) == 1
        check(_lib.lib().vlut_scale_rows(self._h, x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), N, T, scale),
              "vlut_scale_rows")
        return out

    def mul_mat_host(self, w: Weight, x_host: torch.Tensor, out_host: torch.Tensor, w_scale: float = 1.0,
                     algo: int = ALGO_AUTO) -> torch.Tensor:
        """The op on HOST buffers (pinned for asynchronous copies): H2D(x) + op + D2H(out) + sync."""
        assert not x_host.is_cuda and not out_host.is_cuda and x_host.dtype == out_host.dtype == torch.float32
        T, K = x_host.shape
        assert K == w.K and out_host.shape == (T, w.N) and x_host.stride(1) == 1 and out_host.stride(1) == 1
        check(_lib.lib().vlut_mul_mat_host(self._h, w.type, w.K, w.N, T, w.t2.data_ptr(), x_host.data_ptr(),
                                           x_host.stride(0), out_host.data_ptr(), out_host.stride(0), w_scale, algo),
              "vlut_mul_mat_host")
        return out_host

    def measure_int8_peak(self) -> tuple[float, float]:
        """(TOP/s over the device, SM clocks per M128 N256 K32 MMA) of a tcgen05.mma kind::i8 issue loop on every SM."""
        tops, clk = C.c_double(0), C.c_double(0)
        check(_lib.lib().vlut_measure_int8_peak(self._h, C.byref(tops), C.byref(clk)), "vlut_measure_int8_peak")
        return tops.value, clk.value

    def sync(self) -> None:
        check(_lib.lib().vlut_ctx_sync(self._h), "vlut_ctx_sync")

    @property
    def launch_count(self) -> int:
        return int(_lib.lib().vlut_ctx_launch_count(self._h))

    @property
    def last_algo(self) -> int:
        return int(_lib.lib().vlut_ctx_last_algo(self._h))


class Group:
    """vlut_group: ONE process driving G GPUs (output-row tensor parallelism behind the C boundary).  Member 0 is the root:
    activations live there, the gathered result lands in the buffer(s) the caller names."""

    def __init__(self, devices: list[int]):
        if not torch.cuda.is_available():
            raise RuntimeError("vlut_b200 needs CUDA devices (sm_100a); there is no CPU fallback")
        arr = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        check(_lib.lib().vlut_group_create(len(devices), arr, C.byref(h)), "vlut_group_create")
        self._h = h
        self.devices = list(devices)
        self._weights = []

    def rows(self, N: int, member: int) -> tuple[int, int]:
        r0, rows = C.c_int64(0), C.c_int64(0)
        check(_lib.lib().vlut_group_rows(self._h, N, member, C.byref(r0), C.byref(rows)), "vlut_group_rows")
        return r0.value, rows.value

    def upload(self, type_: int, K: int, N: int, gguf: np.ndarray):
        g = np.ascontiguousarray(gguf, dtype=np.uint8)
        assert g.size == nbytes(type_, K, N)
        w = C.c_void_p()
        check(_lib.lib().vlut_group_upload_weights(self._h, type_, K, N, g.ctypes.data, None, C.byref(w)),
              "vlut_group_upload_weights")
        self._weights.append(w)
        return (w, type_, K, N)

    def mul_mat(self, w, x_root: torch.Tensor, outs: list[torch.Tensor], w_scale: float = 1.0) -> None:
        """x on the root; outs = one [T][N] f32 buffer (gather) or one per member (all-gather)."""
        wh, type_, K, N = w
        T = x_root.shape[0]
        assert x_root.dtype == torch.float32 and x_root.stride(1) == 1 and x_root.shape[1] == K
        for o in outs:
            assert o.dtype == torch.float32 and o.shape == (T, N) and o.stride(1) == 1
        arr = (C.c_void_p * len(outs))(*[o.data_ptr() for o in outs])
        check(_lib.lib().vlut_group_mul_mat(self._h, wh, T, x_root.data_ptr(), x_root.stride(0), arr, len(outs),
                                            outs[0].stride(0), w_scale), "vlut_group_mul_mat")

    def sync(self) -> None:
        check(_lib.lib().vlut_group_sync(self._h), "vlut_group_sync")

    def close(self) -> None:
        if getattr(self, "_h", None):
            for w in self._weights:
                _lib.lib().vlut_group_free_weights(self._h, w)
            _lib.lib().vlut_group_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
