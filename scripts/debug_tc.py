import sys, os; sys.path.insert(0, '.')
import numpy as np, torch
import vlut_b200
from vlut_b200 import api
from oracle import oracle as O
ctx = api.Context(0)
K, N = int(os.environ.get("DBG_K", 1024)), int(os.environ.get("DBG_N", 256))
rng = np.random.Generator(np.random.PCG64(1))
w = (rng.integers(0, 3, size=(N, K), dtype=np.int8) - 1).astype(np.float32)
packed = O.Oracle().pack(O.I2_V_4, w)
W = ctx.upload(O.I2_V_4, K, N, packed)
modes = os.environ.get("DBG_MODES", "0,1").split(",")
for ts in modes:
  for fast in ("1", "0"):
    for T in [int(t) for t in os.environ.get("DBG_T", "8,16,32,64,128,200").split(",")]:
        os.environ["VLUT_TC_TS"] = ts; os.environ["VLUT_TC_FAST"] = fast
        x = rng.standard_normal((T, K), dtype=np.float32)
        xd = torch.from_numpy(x).cuda()
        e = ctx.mul_mat(W, xd, algo=api.ALGO_SIMPLE).clone()
        torch.cuda.synchronize()
        r = ctx.mul_mat(W, xd, algo=int(os.environ.get('DBG_ALGO', api.ALGO_TC))).clone()
        torch.cuda.synchronize()
        bad = (r != e)
        msg = ""
        if bad.any():
            idx = bad.nonzero()
            msg = f" rows(t) {sorted(set(idx[:,0].tolist()))[:6]}.. cols {sorted(set(idx[:,1].tolist()))[:10]} got {r[idx[0][0], idx[0][1]].item()} exp {e[idx[0][0], idx[0][1]].item()}"
        print(f"TS={ts} fast={fast} T={T}: bad {int(bad.sum())}/{bad.numel()}{msg}", flush=True)
