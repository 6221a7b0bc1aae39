"""time_quant.py -- the activation quantise kernel alone: us per call and algorithmic GB/s (4TK read + TK written) per (K, T),
graph replays over rotating fp32 inputs (> L2 in total for the large cases).  usage: python scripts/time_quant.py"""
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


import os

os.environ["VLUT_QUANT_KERNEL"] = sys.argv[1] if len(sys.argv) > 1 else "0"
print("VLUT_QUANT_KERNEL =", os.environ["VLUT_QUANT_KERNEL"], "(0 register-resident rows, 2 two-pass)")
ctx = api.Context(dev, stream=stream)
for tn, K in [("I2_V_4", 4096), ("I2_V_4", 14336), ("I2_V_4", 2048), ("I2_V_4", 8192), ("I1_V_2", 3200), ("I1_V_2", 8640)]:
    type_ = api.TYPE_BY_NAME[tn]
    row = f"{tn} K={K}:"
    for T in (1, 16, 64, 128, 256, 512, 1024, 2048):
        nbuf = max(2, min(16, int(300e6 // (4 * T * K))))
        xs = [torch.randn((T, K), device=dev, generator=gen) for _ in range(nbuf)]
        us = timed(lambda i: ctx.quantize_act(type_, xs[i % nbuf]), 20)
        row += f"  T={T} {us:5.1f}us {5.0 * T * K / us / 1e3:5.0f}GB/s"
        del xs
    print(row, flush=True)
ctx.close()
