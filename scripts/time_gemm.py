"""Time the GEMM kernel alone (CUDA events, rotating weights) for a few T; env knobs pass through."""
import sys, os; sys.path.insert(0, '.')
import torch
import vlut_b200
from vlut_b200 import api
K, N = 4096, 14336
ctx = api.Context(0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
ws = []
for i in range(11):
    gg = torch.randint(0, 81, (api.nbytes(42, K, N),), dtype=torch.uint8, device="cuda", generator=g)
    ws.append(ctx.upload(42, K, N, gg))
algo = int(os.environ.get("ALGO", api.ALGO_TC))
res = []
# bring the clocks up: the GPU idles at 120 MHz and needs ~1 s of load to reach 1965 MHz
import time
xw = torch.randn((256, K), device="cuda", generator=g); ow = torch.empty((256, N), device="cuda")
ctx.quantize_act(42, xw)
t_end = time.time() + 2.0
while time.time() < t_end:
    for i in range(50): ctx.gemm(ws[i % 11], 256, out=ow)
    torch.cuda.synchronize()
for T in [int(t) for t in (sys.argv[1] if len(sys.argv) > 1 else "1,16,128,512").split(",")]:
    x = torch.randn((T, K), device="cuda", generator=g)
    out = torch.empty((T, N), device="cuda")
    ctx.quantize_act(42, x)
    for i in range(5): ctx.gemm(ws[i % 11], T, out=out, algo=algo)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(40): ctx.gemm(ws[i % 11], T, out=out, algo=algo)
    e1.record(); torch.cuda.synchronize()
    res.append(f"T={T}: {e0.elapsed_time(e1)/40*1e3:.1f}us")
print(os.environ.get("TAG", ""), " ".join(res), flush=True)
