import os, sys
sys.path.insert(0, ".")
import torch
import vlut_b200
from vlut_b200 import api
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
gen = torch.Generator(device=dev); gen.manual_seed(7)
def timed(fn, iters=10):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        for i in range(iters): fn(i)
    g.replay(); torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / iters * 1e3)
    return sorted(ts)[2]
K, N, T = 4096, 14336, 2048
type_ = api.TYPE_BY_NAME["I2_V_4"]
nb = api.nbytes(type_, K, N)
for grid in (148, 112, 74, 36):
    os.environ["VLUT_TC_GRID"] = str(grid); os.environ["VLUT_TC_WHOLE"] = "1"; os.environ["VLUT_TC_NT"] = "128"
    ctx = api.Context(dev, stream=stream)
    ws = [ctx.upload(type_, K, N, torch.randint(0, 81, (nb,), dtype=torch.uint8, device=dev, generator=gen)) for _ in range(12)]
    x = torch.randn((T, K), device=dev, generator=gen); o = torch.empty((T, N), device=dev)
    ctx.quantize_act(type_, x)
    us = timed(lambda i: ctx.gemm(ws[i % 12], T, out=o, algo=api.ALGO_TCP))
    tiles = 56 * 16 / (grid / 2)
    print(f"grid {grid:3d} CTAs: {us:6.1f} us, {tiles:5.1f} tiles per pair -> {us / tiles * 1965:7.0f} clk per tile (8192 of MMA)", flush=True)
    ctx.close(); del ws; torch.cuda.empty_cache()
