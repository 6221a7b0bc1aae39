import sys, os; sys.path.insert(0, '.')
import numpy as np, torch
import vlut_b200
from vlut_b200 import api
from oracle import oracle as O
ctx = api.Context(0)
type_, K, N = O.I1_V_2, 3200, 1280
rng = np.random.Generator(np.random.PCG64(1))
w = (rng.integers(0, 3, size=(N, K), dtype=np.int8) - 1).astype(np.float32)
packed = O.Oracle().pack(type_, w)
W = ctx.upload(type_, K, N, packed)
for T in (16, 64, 65):
    x = rng.standard_normal((T, K), dtype=np.float32)
    xd = torch.from_numpy(x).cuda()
    e = ctx.mul_mat(W, xd, algo=api.ALGO_SIMPLE).clone()
    torch.cuda.synchronize()
    for fast in ("1", "0"):
        for grid in ("40", "20", "13"):
            os.environ["VLUT_TC_GRID"] = grid; os.environ["VLUT_TC_FAST"] = fast
            res = []
            for rep in range(4):
                r = ctx.mul_mat(W, xd, algo=api.ALGO_TC).clone()
                torch.cuda.synchronize()
                bad = (r != e)
                res.append((int(bad.sum()), sorted(set((bad.nonzero()[:,1]//128).tolist()))[:8]))
            print(f"T={T} fast={fast} grid={grid}: {res}", flush=True)
