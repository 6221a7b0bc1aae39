"""timeline_tcs.py -- in-kernel timeline of one CTA of VLUT_ALGO_TCS (VLUT_TC_DEBUG=1): clock64 marks per role.

usage: python scripts/timeline_tcs.py [T] [N] [split] [cta]
"""
import os
import sys

sys.path.insert(0, ".")
os.environ.setdefault("VLUT_TC_DEBUG", "1")
T = int(sys.argv[1]) if len(sys.argv) > 1 else 16
N = int(sys.argv[2]) if len(sys.argv) > 2 else 14336
if len(sys.argv) > 3:
    os.environ["VLUT_TCS_SPLIT"] = sys.argv[3]
if len(sys.argv) > 4:
    os.environ["VLUT_TCS_DBG_CTA"] = sys.argv[4]

import numpy as np
import torch

import vlut_b200  # noqa: F401
from vlut_b200 import _lib, api

K = 4096
ctx = api.Context(0)
g = torch.Generator(device="cuda")
g.manual_seed(1)
Ws = []
for i in range(10):
    gg = torch.randint(0, 81, (api.nbytes(42, K, N),), dtype=torch.uint8, device="cuda", generator=g)
    Ws.append(ctx.upload(42, K, N, gg))
x = torch.randn((T, K), device="cuda", generator=g)
out = torch.empty((T, N), device="cuda")
for mode in ("gemm-only (activations ready)", "quantise+gemm"):
    for i in range(4):
        if mode.startswith("gemm"):
            ctx.quantize_act(42, x)
            torch.cuda.synchronize()
            ctx.gemm(Ws[i], T, out=out, algo=api.ALGO_TCS)
        else:
            ctx.mul_mat(Ws[i + 4], x, out=out, algo=api.ALGO_TCS)
        torch.cuda.synchronize()
    buf = np.zeros(2048, dtype=np.int64)
    rc = _lib.lib().vlut_debug_tc_counters(buf.ctypes.data, buf.size)
    ev = buf.reshape(4, 256, 2)
    t0 = min([int(ev[r, 0, 0]) for r in range(4) if ev[r, 0, 0] > 0] or [0])
    names = ["decode w4", "mma", "epilogue w0", "act producer"]
    kinds = [{10: "start", 0: "data", 1: "Afree", 2: "st-issued", 3: "st-done"}, {10: "start", 0: "Afull", 1: "issued"},
             {10: "start", 0: "accfull", 1: "drained", 2: "teardown"}, {0: "pdl-passed", 1: "copy"}]
    print(f"== T={T} N={N} {mode} rc={rc}")
    for r in range(4):
        items = [(int(t) - t0, kinds[r].get(int(k), str(int(k)))) for t, k in ev[r] if t > 0]
        print(f"  {names[r]:13s}", " ".join(f"{t}:{k}" for t, k in items))
