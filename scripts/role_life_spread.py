import os, sys
sys.path.insert(0, ".")
os.environ["VLUT_TC_DEBUG"] = "1"; os.environ["VLUT_TC_NT"] = "128"
import numpy as np, torch, time
import vlut_b200
from vlut_b200 import api, _lib
K, N = 4096, 14336
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(1)
ctx = api.Context(0)
ws = [ctx.upload(42, K, N, torch.randint(0, 81, (api.nbytes(42, K, N),), dtype=torch.uint8, device=dev, generator=g)) for _ in range(12)]
for T in (128, 512, 1024):
    x = torch.randn((T, K), device=dev, generator=g); out = torch.empty((T, N), device=dev)
    ctx.quantize_act(42, x)
    t_end = time.time() + 0.5
    while time.time() < t_end:
        for i in range(100): ctx.gemm(ws[i % 12], T, out=out, algo=api.ALGO_TCP)
        torch.cuda.synchronize()
    buf = np.zeros(256 * 4 * 8 + 4 * 24 * 2, dtype=np.int64)
    _lib.lib().vlut_debug_tc_counters(buf.ctypes.data, buf.size)
    c = buf[:8192].reshape(256, 4, 8)
    lead = c[:, 1, 7] > 0
    il, el = c[lead, 1, 7], c[:148, 3, 7]
    mhz = (c[lead, 1, 7] / c[lead, 1, 5] * 1e3).mean()
    print(f"T={T}: issuer life clk min {il.min()} mean {il.mean():.0f} max {il.max()};  epilogue life min {el[el>0].min()} mean {el[el>0].mean():.0f} max {el.max()};  {mhz:.0f} MHz -> max epilogue life {el.max()/mhz:.1f} us")
    if T == 512:
        pairs = np.nonzero(lead)[0] // 2
        ep = c[lead, 3, 7]
        SU, npairs = 16, len(pairs)
        has_split = [((p + 1) * SU) // npairs > (p * SU) // npairs for p in range(npairs)]
        order = np.argsort(-ep)
        print("  slowest pairs (pair: epilogue life, split unit?):", ", ".join(f"{pairs[i]}: {ep[i]} {'S' if has_split[pairs[i]] else '-'}" for i in order[:12]))
        print("  fastest:", ", ".join(f"{pairs[i]}: {ep[i]} {'S' if has_split[pairs[i]] else '-'}" for i in order[-6:]))
        sp = np.array(has_split)
        print(f"  mean epilogue life: pairs with a split unit {ep[sp[pairs]].mean():.0f}, without {ep[~sp[pairs]].mean():.0f}")
        isl = c[lead, 1, 7]
        print(f"  mean issuer life: pairs with a split unit {isl[sp[pairs]].mean():.0f}, without {isl[~sp[pairs]].mean():.0f}")
        for role, name in ((3, "epilogue"), (2, "decode"), (0, "producer")):
            w = c[lead, role, :]
            print(f"  {name} wait counters [0..2], [6] (mean clk): split " + " ".join(f"{w[sp[pairs], k].mean():.0f}" for k in (0, 1, 2, 6)) +
                  "   no split " + " ".join(f"{w[~sp[pairs], k].mean():.0f}" for k in (0, 1, 2, 6)))
