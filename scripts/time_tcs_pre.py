"""time_tcs_pre.py -- the small-T streaming kernel with 1 / 2 / 3 weight-ring slots requested before the activation tile
(VLUT_TCS_PRE): GEMM-only and quantise+GEMM us at 4096 -> 14336, graph replays over rotating weights."""
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


for tn, K, N in [("I2_V_4", 4096, 14336), ("I2_V_4", 4096, 4096), ("I1_V_2", 3200, 8640)]:
    type_ = api.TYPE_BY_NAME[tn]
    nb = api.nbytes(type_, K, N)
    hi = 81 if tn.startswith("I2") else 243
    for pre in (3, 2, 1):
        os.environ["VLUT_TCS_PRE"] = str(pre)
        ctx = api.Context(dev, stream=stream)
        ws = []
        for _ in range(min(40, max(2, int(1.3 * 126e6 / nb) + 1))):
            gg = torch.randint(0, hi, (nb,), dtype=torch.uint8, device=dev, generator=gen)
            ws.append(ctx.upload(type_, K, N, gg))
        row = f"{tn} {K}->{N} pre={pre}:"
        for T in (1, 16, 32, 64):
            xs = [torch.randn((T, K), device=dev, generator=gen) for _ in range(4)]
            o = torch.empty((T, N), device=dev)
            ctx.quantize_act(type_, xs[0])
            g_us = timed(lambda i: ctx.gemm(ws[i % len(ws)], T, out=o))
            op_us = timed(lambda i: ctx.mul_mat(ws[i % len(ws)], xs[i % 4], out=o))
            ctx.set_option(api.Context.OPT_INPUTS_VISIBLE, 1)
            opv_us = timed(lambda i: ctx.mul_mat(ws[i % len(ws)], xs[i % 4], out=o))
            ctx.set_option(api.Context.OPT_INPUTS_VISIBLE, 0)
            row += f"  T={T} gemm {g_us:5.2f} op {op_us:5.2f} op(visible) {opv_us:5.2f}"
        print(row, flush=True)
        ctx.close()
        del ws
        torch.cuda.empty_cache()
