"""e2e timing of vlut_mul_mat_host (pinned host buffers) per token-chunk size: VLUT_HOST_CHUNK=0 is the unpipelined op."""
import os, sys, time; sys.path.insert(0, '.')
import torch
import vlut_b200
from vlut_b200 import api
K, N = 4096, 14336
T = int(sys.argv[1]) if len(sys.argv) > 1 else 512
ctx = api.Context(0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
W = ctx.upload(42, K, N, torch.randint(0, 81, (api.nbytes(42, K, N),), dtype=torch.uint8, device="cuda", generator=g))
xh = torch.randn((T, K)).pin_memory()
oh = torch.empty((T, N)).pin_memory()
for i in range(5): ctx.mul_mat_host(W, xh, oh)
t0 = time.perf_counter()
n = 30
for i in range(n): ctx.mul_mat_host(W, xh, oh)
dt = (time.perf_counter() - t0) / n
print(f"chunk={os.environ.get('VLUT_HOST_CHUNK','128')} T={T}: {dt*1e6:.1f} us/op, {(4*T*K+4*T*N)/dt/1e9:.1f} GB/s PCIe both ways, {2*K*N*T/dt/1e12:.1f} Tint-op/s")
