"""Quick parity + timing check of the CTA-pair kernel (ALGO_TCP) against the scalar anchor (ALGO_SIMPLE) on the GPU."""
import sys, time
sys.path.insert(0, '.')
import torch
import vlut_b200
from vlut_b200 import api

ctx = api.Context(0)
g = torch.Generator(device="cuda"); g.manual_seed(3)
ok = True
for (typ, K, N, T) in [(42, 512, 256, 64), (42, 512, 256, 128), (42, 1024, 300, 100), (42, 4096, 1000, 128), (40, 4096, 14336, 512),
                       (42, 4096, 14336, 200), (45, 2560, 640, 333), (42, 4096, 14336, 2048)]:
    gg = torch.randint(0, 81, (api.nbytes(typ, K, N),), dtype=torch.uint8, device="cuda", generator=g)
    W = ctx.upload(typ, K, N, gg)
    for big in (False, True):
        x = torch.randn((T, K), device="cuda", generator=g)
        if big:
            x = torch.sign(x) * (0.8 + 0.2 * torch.rand_like(x))   # |q| large: exact (wrapping) path
        ref = ctx.mul_mat(W, x, algo=api.ALGO_SIMPLE)
        out = ctx.mul_mat(W, x, algo=api.ALGO_TCP)
        torch.cuda.synchronize()
        same = torch.equal(ref.view(torch.int32), out.view(torch.int32))
        ok &= same
        print(f"type={typ} K={K} N={N} T={T} big={big} bit-exact={same}", flush=True)
        if not same:
            d = (ref != out)
            idx = d.nonzero()
            print("   mismatches", int(d.sum()), "first", idx[:4].tolist(), ref[d][:4].tolist(), out[d][:4].tolist())
print("ALL OK" if ok else "FAILED")
# timing
K, N = 4096, 14336
gg = torch.randint(0, 81, (api.nbytes(42, K, N),), dtype=torch.uint8, device="cuda", generator=g)
W = ctx.upload(42, K, N, gg)
t_end = time.time() + 2.0
x = torch.randn((512, K), device="cuda", generator=g)
out = torch.empty((512, N), device="cuda")
while time.time() < t_end:
    ctx.mul_mat(W, x, out=out)
torch.cuda.synchronize()
for T in (64, 128, 256, 512, 1024, 2048):
    x = torch.randn((T, K), device="cuda", generator=g)
    out = torch.empty((T, N), device="cuda")
    ctx.quantize(42, x) if hasattr(ctx, "quantize") else None
    for name, algo in (("auto", api.ALGO_AUTO), ("tcp", api.ALGO_TCP)):
        for _ in range(5):
            ctx.mul_mat(W, x, out=out, algo=algo)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            ctx.mul_mat(W, x, out=out, algo=algo)
        e1.record(); torch.cuda.synchronize()
        print(f"T={T} {name}: {e0.elapsed_time(e1) / 50 * 1000:.1f} us/op (quantize+gemm)", flush=True)
