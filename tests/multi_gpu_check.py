"""Multi-GPU parity + timing of the row-sharded path (run under torchrun on a multi-GPU box):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/multi_gpu_check.py

Every rank checks that the NCCL all-gather path and the fused (epilogue-scatter) path both reproduce the unsharded
oracle result bit for bit, then times both."""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import vlut_b200  # noqa: E402,F401
from vlut_b200 import api, shard  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = api.Context(local)
    orc = O.Oracle()
    ok = True
    for type_, K, N, T in [(O.I2_V_4, 3072, 9216, 16), (O.I2_V_4, 3072, 3072, 512), (O.I1_V_2, 3200, 8640, 33)]:
        w = O.make_ternary(N, K, seed=1)
        x = O.make_acts(T, K, seed=2)
        packed = orc.pack(type_, w)
        ref = orc.mul_mat(type_, packed, x, N)
        W = shard.RowShardedWeight(ctx, type_, K, N, packed)
        xd = torch.from_numpy(x).cuda()
        out = W.mul_mat(xd)
        torch.cuda.synchronize()
        same = np.array_equal(out.cpu().numpy().view(np.uint32), ref.view(np.uint32))
        msg = f"rank {rank}: {O.TYPE_NAMES[type_]} K={K} N={N} T={T} nccl-gather bit-exact={same}"
        ok &= same
        try:
            outf = W.mul_mat_fused(xd)
            torch.cuda.synchronize()
            samef = np.array_equal(outf.cpu().numpy().view(np.uint32), ref.view(np.uint32))
            _, hdl = W.symm_out(T)
            msg += f" fused bit-exact={samef} (multicast={'yes' if int(getattr(hdl, 'multicast_ptr', 0) or 0) else 'no'})"
            ok &= samef
        except Exception as e:  # noqa: BLE001
            msg += f" fused FAILED: {type(e).__name__}: {e}"
            ok = False
        print(msg, flush=True)
        # repeated calls rotate the result buffers: every one of them must hold the right bits
        try:
            for rep in range(5):
                o = W.mul_mat_fused(xd)
                torch.cuda.synchronize()
                ok &= bool(np.array_equal(o.cpu().numpy().view(np.uint32), ref.view(np.uint32)))
        except Exception as e:  # noqa: BLE001
            print(f"rank {rank}: repeated fused calls FAILED: {e}", flush=True)
            ok = False

        def fused_scalar(xx):
            return W.mul_mat_fused(xx)

        # timing, device events, max over ranks
        for name, fn in (("nccl", W.mul_mat), ("fused (16-byte scatter stores)", W.mul_mat_fused), ("fused (4-byte scatter stores)", fused_scalar)):
            os.environ["VLUT_SCATTER_VEC"] = "0" if "4-byte" in name else "1"
            api.reload_env()
            try:
                for _ in range(5):
                    fn(xd)
                dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(20):
                    fn(xd)
                e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / 20 * 1e3], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                if rank == 0:
                    print(f"   {name}: {t.item():.1f} us per op (max over ranks)", flush=True)
            except Exception as e:  # noqa: BLE001
                if rank == 0:
                    print(f"   {name}: timing failed: {e}", flush=True)
    os.environ["VLUT_SCATTER_VEC"] = "1"
    api.reload_env()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("MULTI_GPU_CHECK", "PASS" if flag.item() == 1 else "FAIL", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
