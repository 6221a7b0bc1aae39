"""Per-role wait-cycle breakdown of the tcgen05 GEMM (VLUT_TC_DEBUG=1)."""
import sys, os; sys.path.insert(0, '.')
os.environ["VLUT_TC_DEBUG"] = "1"

import numpy as np, torch
import vlut_b200
from vlut_b200 import api, _lib
K, N = 4096, 14336
ctx = api.Context(0)
g = torch.Generator(device="cuda"); g.manual_seed(1)
gg = torch.randint(0, 81, (api.nbytes(42, K, N),), dtype=torch.uint8, device="cuda", generator=g)
W = ctx.upload(42, K, N, gg)
for T in [int(t) for t in (sys.argv[1] if len(sys.argv) > 1 else "16,128,512").split(",")]:
    x = torch.randn((T, K), device="cuda", generator=g)
    out = torch.empty((T, N), device="cuda")
    for i in range(3):
        ctx.mul_mat(W, x, out=out)
    torch.cuda.synchronize()
    buf = np.zeros(256 * 4 * 8 + 4 * 24 * 2, dtype=np.int64)
    rc = _lib.lib().vlut_debug_tc_counters(buf.ctypes.data, buf.size)
    c = buf[:8192].reshape(256, 4, 8)
    ev = buf[8192:].reshape(4, 24, 2)
    for name, used in (("leader CTAs", c[:, 1, 7] > 0), ("peer CTAs", (c[:, 1, 7] == 0) & (c[:, 2, 7] > 0))):
        n = int(used.sum())
        if n == 0:
            continue
        m = c[used].mean(axis=0)
        print(f"T={T} rc={rc} {name}: {n}   (mean cycles per CTA; life = role start to role exit)")
        print(f"  act producer: life {m[0,7]:.0f}  wait_empty {m[0,0]:.0f}")
        print(f"  mma issuer  : life {m[1,7]:.0f}  wait_afull {m[1,1]:.0f}  wait_accempty {m[1,2]:.0f}")
        print(f"  decode grp 0: life {m[2,7]:.0f}  wait_full {m[2,0]:.0f}  wait_aempty {m[2,1]:.0f}  wait_st {m[2,2]:.0f}")
        print(f"  decode grp 1:               wait_full {m[2,3]:.0f}  wait_aempty {m[2,4]:.0f}  wait_st {m[2,5]:.0f}")
        print(f"  epilogue    : life {m[3,7]:.0f}  wait_accfull {m[3,0]:.0f}")

    if os.environ.get("TIMELINE"):
        names = ["producer", "mma", "decode0", "epilogue"]
        kinds = {0: "w0", 1: "w1", 2: "st-issued", 3: "fetched", 4: "st-done", 5: "fenced", 6: "chunk", 7: "drained", 8: "mma-issued", 9: "A-ready", 10: "parked", 11: "counted", 12: "fixed-up", 13: "mmas-out", 14: "stage-committed"}
        for r in range(4):
            print("  timeline", names[r], " ".join(f"{int(t)}:{kinds.get(int(k), k)}" for t, k in ev[r] if t > 0))
