"""CPU suite, part 3: the N > 1 path's host logic with the gloo backend, world_size 2 (and 3).

Each rank cuts its row shard out of the GGUF-packed matrix (the C ABI's host-side vlut_shard_rows_host), computes its
output slice -- here with the CPU oracle, in the test only; the product computes it on the GPU -- and the slices are
all-gathered with the same all_gather_rows() the NCCL path uses.  The result must equal the unsharded oracle result
bit for bit, including ragged shapes (N not a multiple of 128 x world)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, type_, K, N, T, result_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import vlut_b200  # noqa: F401
    from vlut_b200 import shard
    orc = O.Oracle()
    w = O.make_ternary(N, K, seed=7)
    x = O.make_acts(T, K, seed=8)
    packed = orc.pack(type_, w)
    plan = shard.RowPlan(N, world)
    row0, rows = plan.rows(rank)
    local_packed = shard.shard_gguf(type_, K, N, plan, rank, packed)
    assert np.array_equal(local_packed, orc.pack(type_, w[row0:row0 + rows]))
    local = torch.from_numpy(orc.mul_mat(type_, local_packed, x, rows, nth=1))
    full = shard.all_gather_rows(local, plan)
    ref = orc.mul_mat(type_, packed, x, N, nth=1)
    ok = np.array_equal(full.numpy().view(np.uint32), ref.view(np.uint32))
    open(os.path.join(result_dir, f"rank{rank}.txt"), "w").write("ok" if ok else "mismatch")
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("type_,K,N,T", [(O.I2_V_4, 1024, 512, 9), (O.I1_V_2, 1300, 300, 5), (O.I2_V, 512, 130, 3)])
def test_row_sharded_all_gather_gloo(tmp_path, world, type_, K, N, T):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, type_, K, N, T, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert (tmp_path / f"rank{r}.txt").read_text() == "ok"


def test_row_plan():
    import vlut_b200  # noqa: F401
    from vlut_b200.shard import RowPlan
    p = RowPlan(14336, 8)
    assert p.bounds() == [i * 1792 for i in range(9)]
    p = RowPlan(8640, 4)  # 67.5 strips: shards of 17,17,17,17 strips, the last one cut at N
    b = p.bounds()
    assert b[0] == 0 and b[-1] == 8640 and all(b[i] < b[i + 1] for i in range(4))
    assert all((b[i + 1] - b[i]) % 128 == 0 for i in range(3))
    p = RowPlan(130, 4)  # fewer strips than ranks: trailing ranks own nothing
    assert p.bounds() == [0, 128, 130, 130, 130]
