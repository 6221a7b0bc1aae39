import os, sys
sys.path.insert(0, ".")
os.environ["VLUT_TC_DEBUG"] = "1"; os.environ["VLUT_TC_NT"] = "128"
import numpy as np, torch, time
import vlut_b200
from vlut_b200 import api, _lib
K, N = 4096, 14336
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(1)
for abl in (0, 8, 16):
    os.environ["VLUT_TC_ABL"] = str(abl)
    ctx = api.Context(0)
    ws = [ctx.upload(42, K, N, torch.randint(0, 81, (api.nbytes(42, K, N),), dtype=torch.uint8, device=dev, generator=g)) for _ in range(12)]
    for T in (512, 2048):
        x = torch.randn((T, K), device=dev, generator=g); out = torch.empty((T, N), device=dev)
        ctx.quantize_act(42, x)
        t_end = time.time() + 0.4
        while time.time() < t_end:
            for i in range(100): ctx.gemm(ws[i % 12], T, out=out, algo=api.ALGO_TCP)
            torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(100): ctx.gemm(ws[i % 12], T, out=out, algo=api.ALGO_TCP)
        e1.record(); torch.cuda.synchronize()
        buf = np.zeros(256 * 4 * 8 + 4 * 24 * 2, dtype=np.int64)
        _lib.lib().vlut_debug_tc_counters(buf.ctypes.data, buf.size)
        c = buf[:8192].reshape(256, 4, 8)
        lead = c[:, 1, 7] > 0
        print(f"abl={abl} T={T}: {e0.elapsed_time(e1)*10:.1f} us/launch (debug build of the launch); issuer life {c[lead,1,7].mean():.0f}, epilogue life {c[lead,3,7].mean():.0f} max {c[lead,3,7].max()}, epilogue accfull wait {c[lead,3,0].mean():.0f} clk")
    ctx.close(); del ws
