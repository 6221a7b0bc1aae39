import sys, os; sys.path.insert(0, '.')
import numpy as np, torch
import vlut_b200
from vlut_b200 import api
from oracle import oracle as O
ctx = api.Context(0)
K, N = 4096, 2048
rng = np.random.Generator(np.random.PCG64(1))
w = (rng.integers(0, 3, size=(N, K), dtype=np.int8) - 1).astype(np.float32)
packed = O.Oracle().pack(O.I2_V_4, w)
W = ctx.upload(O.I2_V_4, K, N, packed)
for smem_kb in (200, 100, 60):
  for grid in (148, 20, 7, 3):
    for T in (8, 16, 17, 32, 48, 64):
        os.environ["VLUT_MMA_SMEM_KB"] = str(smem_kb); os.environ["VLUT_MMA_GRID"] = str(grid)
        x = rng.standard_normal((T, K), dtype=np.float32)
        xd = torch.from_numpy(x).cuda()
        r = ctx.mul_mat(W, xd, algo=api.ALGO_MMA).clone()
        e = ctx.mul_mat(W, xd, algo=api.ALGO_SIMPLE).clone()
        torch.cuda.synchronize()
        bad = (r != e)
        msg = ""
        if bad.any():
            idx = bad.nonzero()
            msg = f" rows {sorted(set(idx[:,0].tolist()))[:6]}.. tiles {sorted(set((idx[:,1]//128).tolist()))[:10]} got {float(r[idx[0][0], idx[0][1]])} exp {float(e[idx[0][0], idx[0][1]])}"
        print(f"smem={smem_kb} grid={grid} T={T}: bad {int(bad.sum())}/{bad.numel()}{msg}")
