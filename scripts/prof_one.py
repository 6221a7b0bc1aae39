"""Run the Vec-LUT MUL_MAT a few times at one token count (profiling target for ncu)."""
import sys, os; sys.path.insert(0, '.')
import torch
import vlut_b200
from vlut_b200 import api
T = int(sys.argv[1]) if len(sys.argv) > 1 else 16
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 6
K, N = [int(v) for v in os.environ.get("VLUT_SHAPE", "4096,14336").split(",")]
type_ = api.TYPE_BY_NAME[os.environ.get("VLUT_TYPE", "I2_V_4")]
ctx = api.Context(0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
ws = []
for i in range(3):
    gg = torch.randint(0, 81 if type_ in (40, 42, 43) else 243, (api.nbytes(type_, K, N),), dtype=torch.uint8, device="cuda", generator=g)
    ws.append(ctx.upload(type_, K, N, gg))
x = torch.randn((T, K), device="cuda", generator=g)
out = torch.empty((T, N), device="cuda")
for i in range(iters):
    ctx.mul_mat(ws[i % 3], x, out=out)
torch.cuda.synchronize()
print("done", T, api.ALGO_NAMES[ctx.last_algo])
