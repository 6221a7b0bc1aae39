"""time_candidates.py -- the small-token contenders side by side (whole op = quantise + GEMM + epilogue, graph replays, weights
rotating over more than L2): TCS (tcgen05 streaming kernel + quantise kernel), GEMV (CUDA cores, quantisation fused, one
launch), LUT (faithful shared-memory vector-LUT).  With PROFILE=<algo>,<T> it runs that contender a few times un-timed (for
`ncu -k regex:...`).

usage: python scripts/time_candidates.py [shape ...]     shape = TYPE:K:N, default I2_V_4:4096:14336 I1_V_2:3200:8640
"""
import os
import sys

sys.path.insert(0, ".")
import torch

import vlut_b200  # noqa: F401
from vlut_b200 import api

L2 = 126 * 1024 * 1024
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(stream)
ctx = api.Context(dev, stream=stream)
gen = torch.Generator(device=dev)
gen.manual_seed(7)
HBM = 6455.9


def weights_for(type_, K, N):
    nb = api.nbytes(type_, K, N)
    n_w = min(40, max(2, -(-int(1.25 * L2) // nb)))
    hi = 81 if type_ in (api.I2_V, api.I2_V_4, api.I2_V_8) else 243
    ws = []
    for _ in range(n_w):
        gg = torch.randint(0, hi, (nb,), dtype=torch.uint8, device=dev, generator=gen)
        ws.append(ctx.upload(type_, K, N, gg))
    return ws, nb


def timed(fn, iters=60):
    for i in range(5):
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


shapes = sys.argv[1:] or ["I2_V_4:4096:14336", "I1_V_2:3200:8640", "I2_V_4:4096:4096", "I2_V_4:14336:4096"]
prof = os.environ.get("PROFILE")
for sh in shapes:
    tn, K, N = sh.split(":")
    K, N = int(K), int(N)
    type_ = api.TYPE_BY_NAME[tn]
    ws, nb = weights_for(type_, K, N)
    if prof:
        an, T = prof.split(",")
        T = int(T)
        algo = {"tcs": api.ALGO_TCS, "gemv": api.ALGO_GEMV, "lut": api.ALGO_LUT, "tcp": api.ALGO_TCP}[an]
        x = torch.randn((T, K), device=dev, generator=gen)
        o = torch.empty((T, N), device=dev)
        for i in range(6):
            ctx.mul_mat(ws[i % len(ws)], x, out=o, algo=algo)
        torch.cuda.synchronize()
        print(f"profiled {an} T={T} {sh}")
        break
    print(f"== {sh}  ({nb / 1e6:.2f} MB packed)")
    for T in (1, 2, 4, 8, 16):
        x = [torch.randn((T, K), device=dev, generator=gen) for _ in range(4)]
        o = torch.empty((T, N), device=dev)
        row = f"T={T:3d}"
        for name, algo in (("tcs", api.ALGO_TCS), ("gemv", api.ALGO_GEMV), ("lut", api.ALGO_LUT)):
            if algo == api.ALGO_GEMV and T > 8:
                continue
            if algo == api.ALGO_LUT and (T > 16 or N * K > 4096 * 14336):
                continue
            try:
                us = timed(lambda i: ctx.mul_mat(ws[i % len(ws)], x[i % 4], out=o, algo=algo), 60 if algo != api.ALGO_LUT else 10)
                alg = nb + 4 * T * K + 4 * T * N
                row += f"  {name} {us:7.2f} us ({alg / us / 1e3 / HBM * 100:4.1f}% HBM)"
            except Exception as e:  # noqa: BLE001
                row += f"  {name} failed ({e})"
        print(row, flush=True)
    del ws
    torch.cuda.empty_cache()
