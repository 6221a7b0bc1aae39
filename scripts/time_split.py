"""time_split.py -- the CTA-pair kernel's work split policies (VLUT_TC_WHOLE: 0 stream-K, 1 whole tiles, 2 hybrid) per T, GEMM-only us."""
import os, sys
sys.path.insert(0, ".")
import torch
import vlut_b200  # noqa: F401
from vlut_b200 import api
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
gen = torch.Generator(device=dev); gen.manual_seed(7)
def timed(fn, iters=20):
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
shapes = [("I2_V_4", 4096, 14336), ("I2_V_4", 14336, 4096), ("I2_V_4", 4096, 4096), ("I1_V_2", 3200, 8640)]
for tn, K, N in shapes:
    type_ = api.TYPE_BY_NAME[tn]
    nb = api.nbytes(type_, K, N)
    for whole in (-1, 0, 1, 2):
        os.environ["VLUT_TC_WHOLE"] = str(whole)
        os.environ["VLUT_TC_NT"] = "128"
        ctx = api.Context(dev, stream=stream)
        ws = [ctx.upload(type_, K, N, torch.randint(0, 81 if tn.startswith("I2") else 243, (nb,), dtype=torch.uint8, device=dev, generator=gen)) for _ in range(12)]
        row = f"{tn} {K}->{N} whole={whole:2d}:"
        for T in (192, 256, 384, 512, 640, 768, 1024):
            x = torch.randn((T, K), device=dev, generator=gen)
            o = torch.empty((T, N), device=dev)
            ctx.quantize_act(type_, x)
            us = timed(lambda i: ctx.gemm(ws[i % 12], T, out=o, algo=api.ALGO_TCP))
            row += f"  T={T} {us:5.1f}"
        print(row, flush=True)
        ctx.close(); del ws; torch.cuda.empty_cache()
