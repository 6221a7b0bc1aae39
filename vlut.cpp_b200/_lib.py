"""ctypes binding of libvlut_b200.so (the C ABI of include/vlut_b200.h).

The library is the product: if it is missing or cannot be loaded, importing the compute API raises -- there is
no PyTorch/NumPy/CPU fallback anywhere in this package.
"""
from __future__ import annotations

import ctypes as C
import re
from pathlib import Path

HERE = Path(__file__).resolve().parent
import os

# VLUT_B200_LIB: an alternative build of the same library (instrumented variants used by scripts/timeline_tcs.py)
LIB_PATH = Path(os.environ.get("VLUT_B200_LIB") or (HERE / "csrc" / "build" / "libvlut_b200.so"))
HEADER = HERE.parent / "include" / "vlut_b200.h"

_lib = None

i64, cint, vp, f32 = C.c_int64, C.c_int, C.c_void_p, C.c_float

_SIGNATURES = {
    "vlut_version": (cint, []),
    "vlut_error_string": (C.c_char_p, [cint]),
    "vlut_last_cuda_error": (C.c_char_p, []),
    "vlut_device_count": (cint, [C.POINTER(cint)]),
    "vlut_device_info": (cint, [cint, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t),
                                C.POINTER(cint)]),
    "vlut_ctx_create": (cint, [cint, vp, C.POINTER(vp)]),
    "vlut_ctx_destroy": (None, [vp]),
    "vlut_ctx_stream": (vp, [vp]),
    "vlut_ctx_sync": (cint, [vp]),
    "vlut_ctx_set_option": (cint, [vp, cint, cint]),
    "vlut_ctx_launch_count": (i64, [vp]),
    "vlut_ctx_last_algo": (cint, [vp]),
    "vlut_dev_malloc": (cint, [cint, C.c_size_t, C.POINTER(vp)]),
    "vlut_dev_free": (cint, [cint, vp]),
    "vlut_dev_memcpy_h2d": (cint, [vp, vp, vp, C.c_size_t]),
    "vlut_dev_memcpy_d2h": (cint, [vp, vp, vp, C.c_size_t]),
    "vlut_dev_memset": (cint, [vp, vp, cint, C.c_size_t]),
    "vlut_gguf_nbytes": (C.c_size_t, [cint, i64, i64]),
    "vlut_repacked_nbytes": (C.c_size_t, [cint, i64, i64]),
    "vlut_repack_weights": (cint, [vp, cint, i64, i64, vp, vp]),
    "vlut_upload_weights": (cint, [vp, cint, i64, i64, vp, vp]),
    "vlut_shard_rows_host": (cint, [cint, i64, i64, i64, i64, vp, vp]),
    "vlut_quantize_act": (cint, [vp, cint, vp, i64, i64, i64]),
    "vlut_gemm": (cint, [vp, cint, i64, i64, i64, vp, vp, i64, f32, cint]),
    "vlut_mul_mat": (cint, [vp, cint, i64, i64, i64, vp, vp, i64, vp, i64, f32, cint]),
    "vlut_mul_mat_scatter": (cint, [vp, cint, i64, i64, i64, vp, vp, i64, vp, cint, cint, i64, f32, cint]),
    "vlut_mul_mat_scatter_push": (cint, [vp, cint, i64, i64, i64, vp, vp, i64, vp, cint, cint, i64, f32]),
    "vlut_mul_mat_no_alloc": (cint, [cint, vp, cint, i64, i64, i64, vp, vp, i64, vp, i64, f32, C.c_char_p, C.c_size_t]),
    "vlut_group_create": (cint, [cint, vp, C.POINTER(vp)]),
    "vlut_group_destroy": (None, [vp]),
    "vlut_group_size": (cint, [vp]),
    "vlut_group_ctx": (vp, [vp, cint]),
    "vlut_group_rows": (cint, [vp, i64, cint, C.POINTER(i64), C.POINTER(i64)]),
    "vlut_group_shard0_nbytes": (C.c_size_t, [vp, cint, i64, i64]),
    "vlut_group_upload_weights": (cint, [vp, cint, i64, i64, vp, vp, C.POINTER(vp)]),
    "vlut_group_free_weights": (None, [vp, vp]),
    "vlut_group_mul_mat": (cint, [vp, vp, i64, vp, i64, vp, cint, i64, f32]),
    "vlut_group_sync": (cint, [vp]),
    "vlut_barrier": (cint, [vp, vp, cint, cint]),
    "vlut_mul_mat_host": (cint, [vp, cint, i64, i64, i64, vp, vp, i64, vp, i64, f32, cint]),
    "vlut_scale_rows": (cint, [vp, vp, i64, vp, i64, i64, i64, f32]),
    "vlut_pick_algo": (cint, [cint, i64, i64, i64]),
    "vlut_op_supported": (cint, [cint, i64, i64, i64]),
    "vlut_reload_env": (cint, []),
    "vlut_measure_int8_peak": (cint, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "vlut_mul_mat_batch": (cint, [vp, cint, vp, i64, vp, i64, vp, vp, i64, vp, vp, vp, cint]),
    "vlut_gemm_batch": (cint, [vp, cint, vp, i64, vp, i64, vp, vp, vp, vp, cint]),
    "vlut_debug_get_act": (cint, [vp, i64, i64, vp, vp]),
    "vlut_debug_unpack_t2": (cint, [vp, cint, i64, i64, vp, vp]),
    "vlut_debug_tc_counters": (cint, [vp, cint]),
}


def declared_symbols() -> list[str]:
    """Every function include/vlut_b200.h declares (used by the CPU-side export test)."""
    text = HEADER.read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vlut_[a-z0-9_]+)\s*\(", text)))


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                               f"g.build()'` (nvcc, sm_100a). There is no CPU fallback.")
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class VlutError(RuntimeError):
    def __init__(self, code: int, where: str):
        L = lib()
        msg = L.vlut_error_string(code).decode()
        if code == -2:
            msg += f" [{L.vlut_last_cuda_error().decode()}]"
        super().__init__(f"{where}: {msg} (code {code})")
        self.code = code


def check(code: int, where: str) -> None:
    if code != 0:
        raise VlutError(code, where)
