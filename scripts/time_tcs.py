"""time_tcs.py -- GPU time (CUDA graph replay) of VLUT_ALGO_TCS: GEMM only and quantise+GEMM, a few shapes; used to compare builds."""
import os, sys
sys.path.insert(0, ".")
import torch
import vlut_b200  # noqa: F401
from vlut_b200 import api
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
ctx = api.Context(dev, stream=stream)
gen = torch.Generator(device=dev); gen.manual_seed(7)
def timed(fn, iters=100):
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
algo = int(os.environ.get("ALGO", api.ALGO_TCS))
shapes = [(42, 4096, 14336), (42, 4096, 4096)] if len(sys.argv) < 2 else [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
for type_, K, N in shapes:
    nb = api.nbytes(type_, K, N)
    n_w = min(40, max(2, -(-int(1.25 * 126 * 2**20) // nb)))
    hi = 81 if type_ in (api.I2_V, api.I2_V_4, api.I2_V_8) else 243
    ws = [ctx.upload(type_, K, N, torch.randint(0, hi, (nb,), dtype=torch.uint8, device=dev, generator=gen)) for _ in range(n_w)]
    line = []
    for T in [int(t) for t in os.environ.get("TS", "1,16,32,64").split(",")]:
        xs = [torch.randn((T, K), dtype=torch.float32, device=dev, generator=gen) for _ in range(4)]
        o = torch.empty((T, N), dtype=torch.float32, device=dev)
        ctx.quantize_act(type_, xs[0])
        g_us = timed(lambda i: ctx.gemm(ws[i % n_w], T, out=o, algo=algo))
        op_us = timed(lambda i: ctx.mul_mat(ws[i % n_w], xs[i % 4], out=o, algo=algo))
        line.append(f"T={T}: {g_us:.2f}/{op_us:.2f}")
    print(f"{os.environ.get('TAG','')} type={type_} K={K} N={N} gemm/op us  " + "  ".join(line), flush=True)
    del ws; torch.cuda.empty_cache()
