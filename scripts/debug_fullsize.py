import sys; sys.path.insert(0, '.')
import numpy as np, torch
import vlut_b200
from vlut_b200 import api
from oracle import oracle as O
ctx = api.Context(0)
K, N = 4096, 14336
for T in (4, 16, 48):
    rng = np.random.Generator(np.random.PCG64(T))
    w = (rng.integers(0, 3, size=(N, K), dtype=np.int8) - 1).astype(np.float32)
    packed = O.Oracle().pack(O.I2_V_4, w)
    x = rng.standard_normal((T, K), dtype=np.float32)
    W = ctx.upload(O.I2_V_4, K, N, packed)
    xd = torch.from_numpy(x).cuda()
    q = None
    res = {}
    for name, algo in (("mma", api.ALGO_MMA), ("simple", api.ALGO_SIMPLE)):
        res[name] = ctx.mul_mat(W, xd, algo=algo).clone()
    q, s = ctx.get_act(K, T)
    acc = (torch.from_numpy(q).cuda().double() @ torch.from_numpy(w).cuda().double().T)
    expect = (acc.float() * torch.from_numpy(s).cuda()[:, None])
    for name, r in res.items():
        bad = (r != expect)
        print(f"T={T} {name}: mismatches {int(bad.sum())} / {bad.numel()}")
        if bad.any():
            idx = bad.nonzero()
            print("   first:", idx[:8].tolist(), "rows(t) affected:", sorted(set(idx[:,0].tolist()))[:20])
            cols = idx[:,1]
            print("   col tiles affected:", sorted(set((cols//128).tolist()))[:30])
            i0 = idx[0]
            print("   got", float(r[i0[0], i0[1]]), "expect", float(expect[i0[0], i0[1]]), "scale", float(s[i0[0]]))
