"""time_op.py -- GPU time (CUDA graph replay) of quantise+GEMM through AUTO at a list of T, K=4096 N=14336 I2_V_4."""
import os, sys
sys.path.insert(0, ".")
import torch
import vlut_b200  # noqa: F401
from vlut_b200 import api
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = api.Context(dev, stream=stream)
gen = torch.Generator(device=dev); gen.manual_seed(7)
def timed(fn, iters=60):
    for i in range(5): fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        for i in range(iters): fn(i)
    g.replay(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / iters * 1e3)
    return best
type_, K, N = [int(v) for v in os.environ.get("SHAPE", "42,4096,14336").split(",")]
nb = api.nbytes(type_, K, N)
ws = [ctx.upload(type_, K, N, torch.randint(0, 81 if type_ in (40, 42, 43) else 243, (nb,), dtype=torch.uint8, device=dev, generator=gen)) for _ in range(12)]
line = []
for T in [int(t) for t in (sys.argv[1] if len(sys.argv) > 1 else "16,128,512,2048").split(",")]:
    xs = [torch.randn((T, K), dtype=torch.float32, device=dev, generator=gen) for _ in range(4)]
    o = torch.empty((T, N), dtype=torch.float32, device=dev)
    ctx.quantize_act(type_, xs[0])
    g_us = timed(lambda i: ctx.gemm(ws[i % 12], T, out=o))
    q_us = timed(lambda i: ctx.quantize_act(type_, xs[i % 4]))
    op_us = timed(lambda i: ctx.mul_mat(ws[i % 12], xs[i % 4], out=o))
    line.append(f"T={T}: gemm {g_us:.2f} quant {q_us:.2f} op {op_us:.2f}")
print(os.environ.get("TAG", ""), "  ".join(line), flush=True)
