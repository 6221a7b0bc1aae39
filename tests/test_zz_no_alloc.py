"""vlut_mul_mat_no_alloc (include/vlut_b200.h): the entry point that allocates nothing -- its scratch is a file-scope
__device__ array (vlut.cpp_b200/csrc/no_alloc.cu).  Called through the C ABI and compared, bit for bit, with vlut_mul_mat (the
existing entry point) and with the oracle, on two calls in a row that share the scratch.

Collected last on purpose.  Passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def no_alloc(w, x, out, w_scale=1.0):
    from vlut_b200 import _lib
    err = C.create_string_buffer(512)
    T = x.shape[0]
    rc = _lib.lib().vlut_mul_mat_no_alloc(x.device.index, torch.cuda.current_stream().cuda_stream, w.type, w.K, w.N, T,
                                          w.t2.data_ptr(), x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), w_scale,
                                          err, len(err))
    return rc, err.value.decode()


# the second call of every pair is SMALLER in T and differs in K: it finds the first call's rows (and K padding) in the scratch
PAIRS = [
    ((O.I2_V_4, 4096, 300, 37), (O.I1_V_2, 3200, 128, 5)),      # K = 3200: five 640-k groups
    ((O.I1_V, 1300, 77, 9), (O.I2_V, 1024, 130, 3)),            # ragged N, K padded from 1300 to 1408
    ((O.I2_V_8, 2048, 64, 64), (O.I2_V_4, 4096, 256, 64)),
]


@pytest.mark.parametrize("first,second", PAIRS)
def test_no_alloc_matches_mul_mat_on_two_calls_in_a_row(ctx, orc, first, second):
    from vlut_b200 import api
    cases = []
    for i, (type_, K, N, T) in enumerate((first, second)):
        wt = O.make_ternary(N, K, seed=3 + i)
        x = O.make_acts(T, K, seed=17 + i, kind="outlier" if i else "normal")
        packed = orc.pack(type_, wt)
        w = ctx.upload(type_, K, N, packed)
        xd = torch.from_numpy(x).cuda()
        ref = ctx.mul_mat(w, xd, algo=api.ALGO_AUTO)              # the existing entry point
        out = torch.full((T, N), float("nan"), device="cuda")     # allocated by the CALLER, before the calls
        cases.append((type_, packed, x, w, xd, ref, out))
    torch.cuda.synchronize()
    # two calls in a row on one stream, no synchronisation in between
    for _, _, _, w, xd, _, out in cases:
        rc, msg = no_alloc(w, xd, out)
        assert rc == 0, msg
    torch.cuda.synchronize()
    for type_, packed, x, w, xd, ref, out in cases:
        assert np.array_equal(bits(out.cpu().numpy()), bits(ref.cpu().numpy()))
        assert np.array_equal(bits(out.cpu().numpy()), bits(orc.mul_mat(type_, packed, x, w.N)))
    # ... and again in the other order, with a weight scale folded in, device memory untouched by the calls
    for type_, packed, x, w, xd, ref, out in reversed(cases):
        out.fill_(float("nan"))
        ref2 = ctx.mul_mat(w, xd, w_scale=0.37)
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info()[0]
        rc, msg = no_alloc(w, xd, out, w_scale=0.37)
        assert rc == 0, msg
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free1
        assert np.array_equal(bits(out.cpu().numpy()), bits(ref2.cpu().numpy()))


def test_no_alloc_strided_buffers_and_full_scratch(ctx, orc):
    """ldx > K, ldo > N, and the largest token count that fits the 4 MiB scratch at K = 4096 (1023)."""
    from vlut_b200 import api
    type_, K, N, T = O.I2_V_4, 4096, 128, 1023
    g = torch.Generator(device="cuda").manual_seed(5)
    w = ctx.upload(type_, K, N, orc.pack(type_, O.make_ternary(N, K, seed=8)))
    xbuf = torch.randn((T, K + 8), device="cuda", generator=g)
    x = xbuf[:, :K]
    ref = ctx.mul_mat(w, x)
    obuf = torch.zeros((T, N + 4), device="cuda")
    out = obuf[:, :N]
    rc, msg = no_alloc(w, x, out)
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert torch.equal(out.view(torch.int32), ref.view(torch.int32))
    assert float(obuf[:, N:].abs().max()) == 0.0   # nothing written past N


def test_no_alloc_rejects_what_does_not_fit(ctx, orc):
    from vlut_b200 import api
    type_, K, N = O.I2_V_4, 4096, 128
    g = torch.Generator(device="cuda").manual_seed(6)
    w = ctx.upload(type_, K, N, orc.pack(type_, O.make_ternary(N, K, seed=9)))
    T = 1024   # 1024 * 4096 + 4096 bytes > 4 MiB
    x = torch.randn((T, K), device="cuda", generator=g)
    out = torch.full((T, N), 7.0, device="cuda")
    rc, msg = no_alloc(w, x, out)
    assert rc == -3 and "VLUT_NO_ALLOC_MAX_SCRATCH_BYTES" in msg and "4194304" in msg
    torch.cuda.synchronize()
    assert float(out.min()) == 7.0 and float(out.max()) == 7.0   # nothing was launched
    # bad arguments: clear errors, no launch
    from vlut_b200 import _lib
    err = C.create_string_buffer(256)
    rc = _lib.lib().vlut_mul_mat_no_alloc(0, None, 99, K, N, 4, w.t2.data_ptr(), x.data_ptr(), K, out.data_ptr(), N, 1.0, err, 256)
    assert rc == -1 and b"bad type" in err.value
    rc = _lib.lib().vlut_mul_mat_no_alloc(0, None, type_, K, N, 4, None, x.data_ptr(), K, out.data_ptr(), N, 1.0, None, 0)
    assert rc == -1
    torch.cuda.synchronize()
