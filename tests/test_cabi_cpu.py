"""CPU suite, part 2: the C-ABI library loads without a GPU, exports every symbol include/vlut_b200.h declares,
its host-only helpers are right, and compute entry points fail loudly (no CPU fallback)."""
import ctypes as C

import numpy as np
import pytest

import vlut_b200  # noqa: F401
from vlut_b200 import _lib
from oracle import oracle as O


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    declared = _lib.declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/vlut_b200.h but not exported"
    assert set(_lib._SIGNATURES) == set(declared)
    assert L.vlut_version() == 1


@pytest.mark.parametrize("type_,K,N", [(O.I2_V_4, 4096, 14336), (O.I1_V_2, 3200, 8640), (O.I1_V, 4096, 4096),
                                       (O.I2_V, 8640, 3200), (O.I2_V_8, 2048, 8192)])
def test_nbytes_matches_ggml(type_, K, N):
    L = _lib.lib()
    assert L.vlut_gguf_nbytes(type_, K, N) == O.row_size(type_, K) * N
    # T2: 2 bits per weight on the 128x128-padded matrix
    Kp, Np = -(-K // 128) * 128, -(-N // 128) * 128
    assert L.vlut_repacked_nbytes(type_, K, N) == Kp * Np // 4


def test_invalid_shapes_rejected():
    L = _lib.lib()
    assert L.vlut_gguf_nbytes(41, 4096, 16) == 0      # I2_V_2 has no packer in the reference
    assert L.vlut_gguf_nbytes(O.I2_V_4, 4098, 16) == 0  # K % 4
    assert L.vlut_gguf_nbytes(O.I2_V_4, 0, 16) == 0


@pytest.mark.parametrize("type_,K", [(O.I2_V, 1024), (O.I2_V_4, 1024), (O.I2_V_8, 1024), (O.I1_V, 1300),
                                     (O.I1_V_2, 1300), (O.I1_V_2, 1048)])
def test_shard_rows_host(orc, type_, K):
    """Row shard of a packed matrix == packing the row subset (host-side byte shuffle of the C ABI)."""
    from vlut_b200 import api  # needs torch, not a GPU
    N = 96
    w = O.make_ternary(N, K, seed=3)
    packed = orc.pack(type_, w)
    for row0, rows in [(0, 48), (48, 48), (32, 16), (0, 96), (95, 1)]:
        shard = api.shard_rows(type_, K, N, row0, rows, packed)
        assert np.array_equal(shard, orc.pack(type_, w[row0:row0 + rows]))


def test_compute_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    L = _lib.lib()
    h = C.c_void_p()
    rc = L.vlut_ctx_create(0, None, C.byref(h))
    assert rc == -2 and h.value is None
    assert L.vlut_last_cuda_error()
    from vlut_b200 import api
    with pytest.raises(RuntimeError):
        api.Context(0)
