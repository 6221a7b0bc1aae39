"""time_tcp.py -- the CTA-pair kernel (VLUT_ALGO_TCP) per token tile: GEMM-only us at 4096 -> 14336 (and other shapes) with the
128-token tile vs the 256-token tile (VLUT_TC_NT), graph replays over rotating weights.  usage: python scripts/time_tcp.py"""
import os
import sys

sys.path.insert(0, ".")
import torch

import vlut_b200  # noqa: F401
from vlut_b200 import api

dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
gen = torch.Generator(device=dev)
gen.manual_seed(7)
PEAK = 4567.0


def timed(fn, iters):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        for i in range(iters):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / iters * 1e3)
    return sorted(ts)[len(ts) // 2]


for tn, K, N in [("I2_V_4", 4096, 14336), ("I2_V_4", 14336, 4096), ("I2_V_4", 4096, 4096), ("I1_V_2", 3200, 8640)]:
    type_ = api.TYPE_BY_NAME[tn]
    nb = api.nbytes(type_, K, N)
    hi = 81 if tn.startswith("I2") else 243
    for nt in (128, 256):
        os.environ["VLUT_TC_NT"] = str(nt)
        ctx = api.Context(dev, stream=stream)   # re-reads the knobs
        ws = []
        for _ in range(12):
            gg = torch.randint(0, hi, (nb,), dtype=torch.uint8, device=dev, generator=gen)
            ws.append(ctx.upload(type_, K, N, gg))
        row = f"{tn} {K}->{N} NT={nt}:"
        for T in (256, 384, 512, 768, 1024, 2048):
            x = torch.randn((T, K), device=dev, generator=gen)
            o = torch.empty((T, N), device=dev)
            ctx.quantize_act(type_, x)   # AUTO -> TCP for T > 64
            us = timed(lambda i: ctx.gemm(ws[i % 12], T, out=o, algo=api.ALGO_TCP), 20)
            row += f"  T={T} {us:6.1f}us {2.0 * K * N * T / us / 1e6 / PEAK * 100:4.1f}%"
        print(row, flush=True)
        ctx.close()
        del ws
        torch.cuda.empty_cache()
