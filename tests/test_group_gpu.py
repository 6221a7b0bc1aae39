"""vlut_group: one process, G GPUs, row-sharded weights, gather fused into the GEMM epilogues -- the multi-GPU path behind
the C boundary (what the ggml plugin uses for VLUT_B200_GPUS=G).  On a single-GPU box the members all live on GPU 0 (the
API allows it): same code path, peer stores become local stores.  On a multi-GPU box the real thing runs too."""
import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def group_cases():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    cases = [[0, 0], [0, 0, 0]]
    if n >= 2:
        cases.append([0, 1])
    if n >= 4:
        cases.append([0, 1, 2, 3])
    if n >= 8:
        cases.append(list(range(8)))
    return cases


@pytest.mark.parametrize("devices", group_cases(), ids=lambda d: "gpus-" + "-".join(map(str, d)))
@pytest.mark.parametrize("type_,K,N,T", [(O.I2_V_4, 3072, 1152, 16), (O.I2_V_4, 2048, 1000, 300), (O.I1_V_2, 3200, 520, 40),
                                         (O.I2_V, 1040, 129, 3)])
def test_group_gather_and_allgather(orc, devices, type_, K, N, T):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import vlut_b200  # noqa: F401
    from vlut_b200 import api
    grp = api.Group(devices)
    w = O.make_ternary(N, K, seed=N)
    x = O.make_acts(T, K, seed=T)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    W = grp.upload(type_, K, N, packed)
    # shard plan: whole 128-row strips, everything covered exactly once
    b = 0
    for m in range(len(devices)):
        r0, rows = grp.rows(N, m)
        assert r0 == b and (rows % 128 == 0 or r0 + rows == N)
        b += rows
    assert b == N
    xd = torch.from_numpy(x).to(f"cuda:{devices[0]}")
    # gather into one buffer on the root
    out = torch.full((T, N), float("nan"), device=f"cuda:{devices[0]}")
    for rep in range(2):
        grp.mul_mat(W, xd, [out], w_scale=1.0)
    grp.sync()
    assert np.array_equal(bits(out.cpu().numpy()), bits(ref))
    # all-gather: a buffer on every member
    outs = [torch.full((T, N), float("nan"), device=f"cuda:{d}") for d in devices]
    grp.mul_mat(W, xd, outs, w_scale=0.5)
    grp.sync()
    for o in outs:
        assert np.array_equal(bits(o.cpu().numpy()), bits(ref * np.float32(0.5)))
    grp.close()
