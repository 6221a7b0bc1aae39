"""sm_clock_under_load.py -- the SM frequency INSIDE the CTA-pair kernel: SM cycles (clock64) over wall time (globaltimer) of the
MMA issuer's life, from the debug counters (VLUT_TC_DEBUG=1), after `n` back-to-back launches that keep the chip loaded.
nvidia-smi samples every 200 ms and cannot see a 90 us kernel."""
import os, sys
sys.path.insert(0, ".")
os.environ["VLUT_TC_DEBUG"] = "1"
import numpy as np, torch
import vlut_b200
from vlut_b200 import api, _lib
K, N = 4096, 14336
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(1)
for grid in (0, 74):
    if grid: os.environ["VLUT_TC_GRID"] = str(grid); os.environ["VLUT_TC_WHOLE"] = "1"
    os.environ["VLUT_TC_NT"] = "128"
    ctx = api.Context(0)
    ws = [ctx.upload(42, K, N, torch.randint(0, 81, (api.nbytes(42, K, N),), dtype=torch.uint8, device=dev, generator=g)) for _ in range(12)]
    for T in (512, 2048):
        x = torch.randn((T, K), device=dev, generator=g); out = torch.empty((T, N), device=dev)
        ctx.quantize_act(42, x)
        import time
        for n in (3, 300, 3000):
            t_end = time.time() + (1.5 if n == 3000 else 0.0)   # the last round: 1.5 s of load first (clock ramp), then n launches
            while time.time() < t_end:
                for i in range(100): ctx.gemm(ws[i % 12], T, out=out, algo=api.ALGO_TCP)
                torch.cuda.synchronize()
            for i in range(n): ctx.gemm(ws[i % 12], T, out=out, algo=api.ALGO_TCP)
            torch.cuda.synchronize()
            buf = np.zeros(256 * 4 * 8 + 4 * 24 * 2, dtype=np.int64)
            _lib.lib().vlut_debug_tc_counters(buf.ctypes.data, buf.size)
            c = buf[:8192].reshape(256, 4, 8)
            used = c[:, 1, 7] > 0
            life, ns = c[used, 1, 7].mean(), c[used, 1, 5].mean()
            print(f"grid {grid or 148:3d} CTAs T={T:4d} after {n:3d} launches: issuer life {life:9.0f} clk in {ns / 1e3:7.1f} us -> {life / ns * 1e3:6.0f} MHz", flush=True)
    ctx.close(); del ws; torch.cuda.empty_cache()
