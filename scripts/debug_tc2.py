import sys, os; sys.path.insert(0, '.')
import numpy as np, torch
import vlut_b200
from vlut_b200 import api
from oracle import oracle as O
ctx = api.Context(0)
cases = [(O.I1_V_2, 3200, 8640, 65), (O.I1_V_2, 3200, 1280, 65), (O.I2_V_4, 4096, 14336, 512), (O.I2_V_4, 4096, 2048, 512)]
for type_, K, N, T in cases:
    rng = np.random.Generator(np.random.PCG64(1))
    w = (rng.integers(0, 3, size=(N, K), dtype=np.int8) - 1).astype(np.float32)
    packed = O.Oracle().pack(type_, w)
    W = ctx.upload(type_, K, N, packed)
    x = rng.standard_normal((T, K), dtype=np.float32)
    xd = torch.from_numpy(x).cuda()
    e = ctx.mul_mat(W, xd, algo=api.ALGO_TC1).clone()
    torch.cuda.synchronize()
    for grid in os.environ.get("GRIDS", "148,40,7").split(","):
        os.environ["VLUT_TC_GRID"] = grid
        print(f"{O.TYPE_NAMES[type_]} K={K} N={N} T={T} grid={grid}: launching", flush=True)
        r = ctx.mul_mat(W, xd, algo=api.ALGO_TC).clone()
        torch.cuda.synchronize()
        bad = (r != e)
        msg = ""
        if bad.any():
            idx = bad.nonzero()
            msg = f" rows(t) {sorted(set(idx[:,0].tolist()))[:6]}.. tiles {sorted(set((idx[:,1]//128).tolist()))[:12]} got {r[idx[0][0], idx[0][1]].item()} exp {e[idx[0][0], idx[0][1]].item()}"
        print(f"   bad {int(bad.sum())}/{bad.numel()}{msg}", flush=True)
