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
K, N = 4096, 14336
type_ = api.TYPE_BY_NAME["I2_V_4"]; nb = api.nbytes(type_, K, N)
os.environ["VLUT_TC_NT"] = "256"
for env in ({}, {"VLUT_TC_WHOLE": "1"}, {"VLUT_TC_WHOLE": "2"}, {"VLUT_TC_LEAN": "0"}, {"VLUT_TC_ASTAGES": "2"}):
    for k in ("VLUT_TC_WHOLE", "VLUT_TC_LEAN", "VLUT_TC_ASTAGES"): os.environ.pop(k, None)
    os.environ.update(env)
    ctx = api.Context(dev, stream=stream)
    ws = [ctx.upload(type_, K, N, torch.randint(0, 81, (nb,), dtype=torch.uint8, device=dev, generator=gen)) for _ in range(12)]
    row = f"{env}:"
    for T in (768, 1024, 2048):
        x = torch.randn((T, K), device=dev, generator=gen); o = torch.empty((T, N), device=dev)
        ctx.quantize_act(type_, x)
        row += f"  T={T} {timed(lambda i: ctx.gemm(ws[i % 12], T, out=o, algo=api.ALGO_TCP)):6.1f}"
    print(row, flush=True)
    ctx.close(); del ws; torch.cuda.empty_cache()
