"""time_algo.py T[,T..] algo [type]: GEMM-only time (quantize once, then vlut_gemm repeatedly) for one algorithm."""
import sys, time
sys.path.insert(0, '.')
import torch
import vlut_b200
from vlut_b200 import api, _lib

Ts = [int(t) for t in sys.argv[1].split(",")]
algo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
typ = int(sys.argv[3]) if len(sys.argv) > 3 else 42
K, N = 4096, 14336
ctx = api.Context(0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
gg = torch.randint(0, 81, (api.nbytes(typ, K, N),), dtype=torch.uint8, device="cuda", generator=g)
W = ctx.upload(typ, K, N, gg)
x = torch.randn((512, K), device="cuda", generator=g)
out = torch.empty((512, N), device="cuda")
t_end = time.time() + 2.0
while time.time() < t_end:
    ctx.mul_mat(W, x, out=out)
torch.cuda.synchronize()
for T in Ts:
    x = torch.randn((T, K), device="cuda", generator=g)
    out = torch.empty((T, N), device="cuda")
    ctx.quantize_act(typ, x)
    for _ in range(5):
        ctx.gemm(W, T, out=out, algo=algo)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        ctx.gemm(W, T, out=out, algo=algo)
    e1.record(); torch.cuda.synchronize()
    print(f"T={T} algo={algo}: gemm {e0.elapsed_time(e1) / 50 * 1000:.1f} us", flush=True)
