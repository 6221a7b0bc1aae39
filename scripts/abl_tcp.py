"""abl_tcp.py -- where the CTA-pair kernel's time goes: GEMM-only us with parts of the kernel switched off (VLUT_TC_ABL bits:
1 no tcgen05.st, 2 no MMA, 4 no wait::st, 16 no fast-path drain, 32 no activation copies, 64 no weight copies; results are
wrong by design, only the time is read).  usage: python scripts/abl_tcp.py [NT ...]"""
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
K, N = 4096, 14336
type_ = api.TYPE_BY_NAME["I2_V_4"]


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


ABLS = [(0, "full kernel"), (2, "no MMA"), (5, "no tcgen05.st / wait::st"), (16, "no drain"), (32, "no activation copies"),
        (64, "no weight copies"), (96, "no copies at all"), (2 | 5 | 16, "copies + barriers only"),
        (2 | 5 | 16 | 96, "barriers only"), (5 | 16 | 96, "MMA + barriers only"), (5 | 16, "MMA + copies"),
        (5 | 96, "MMA + drain"), (16 | 96, "MMA + tcgen05.st"), (119 | 256, "barriers only, no LDS/unpack"),
        (119 | 256 | 512, "A ring only (no stage ring)"), (117 | 256 | 512, "A ring + MMA only")]
nts = [int(a) for a in sys.argv[1:]] or [128, 256]
nb = api.nbytes(type_, K, N)
for nt in nts:
    os.environ["VLUT_TC_NT"] = str(nt)
    for abl, name in ABLS:
        os.environ["VLUT_TC_ABL"] = str(abl)
        ctx = api.Context(dev, stream=stream)   # re-reads the knobs
        ws = []
        for _ in range(12):
            gg = torch.randint(0, 81, (nb,), dtype=torch.uint8, device=dev, generator=gen)
            ws.append(ctx.upload(type_, K, N, gg))
        row = f"NT={nt} abl={abl:3d} {name:28s}"
        for T in (512, 2048):
            x = torch.randn((T, K), device=dev, generator=gen)
            o = torch.empty((T, N), device=dev)
            ctx.quantize_act(type_, x)
            us = timed(lambda i: ctx.gemm(ws[i % 12], T, out=o, algo=api.ALGO_TCP), 20)
            row += f"  T={T} {us:6.1f}us"
        print(row, flush=True)
        ctx.close()
        del ws
        torch.cuda.empty_cache()
