"""time_coop.py -- hybrid split tiles with at most one split unit per pair: the contributors share the fix-up (VLUT_TC_COOP=1)
against the last arriver doing all of it (0).  GEMM-only us; one CUDA graph of 20 launches (12 weight copies) per setting, the
two graphs replayed ALTERNATELY 9 times each (same process, same thermal state), medians."""
import os, sys
sys.path.insert(0, ".")
import torch
import vlut_b200  # noqa: F401
from vlut_b200 import api
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev); torch.cuda.set_stream(stream)
gen = torch.Generator(device=dev); gen.manual_seed(7)
KNOB = os.environ.get("AB_KNOB", "VLUT_TC_COOP")
VALS = [int(v) for v in os.environ.get("AB_VALS", "0,1").split(",")]   # e.g. AB_KNOB=VLUT_TC_NT AB_VALS=128,256
if os.environ.get("AB_TS"): pass
def capture(fn, iters=20):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        for i in range(iters): fn(i)
    g.replay(); torch.cuda.synchronize()
    return g
def replay(g, iters=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
shapes = [("I2_V_4", 4096, 14336), ("I2_V_4", 14336, 4096), ("I2_V_4", 4096, 4096), ("I1_V_2", 3200, 8640), ("I1_V_2", 8640, 3200)]
Ts = tuple(int(t) for t in os.environ.get("AB_TS", "384,512,640,768,1024,1536,2048").split(","))
ctxs = {v: api.Context(dev, stream=stream) for v in VALS}   # one activation workspace per setting (the layout depends on the tile)
ctx = ctxs[VALS[0]]
for tn, K, N in shapes:
    type_ = api.TYPE_BY_NAME[tn]
    nb = api.nbytes(type_, K, N)
    ws = [ctx.upload(type_, K, N, torch.randint(0, 81, (nb,), dtype=torch.uint8, device=dev, generator=gen)) for _ in range(12)]
    rows = {v: f"{tn} {K}->{N} {KNOB}={v}" for v in VALS}
    for T in Ts:
        x = torch.randn((T, K), device=dev, generator=gen)
        o = torch.empty((T, N), device=dev)
        gs = {}
        for v in VALS:
            os.environ[KNOB] = str(v)
            api.reload_env()
            ctxs[v].quantize_act(type_, x)
            gs[v] = capture(lambda i: ctxs[v].gemm(ws[i % 12], T, out=o, algo=api.ALGO_TCP))
        ts = {v: [] for v in VALS}
        for _ in range(9):
            for v in VALS: ts[v].append(replay(gs[v]))
        for v in VALS: rows[v] += f" {T}:{sorted(ts[v])[4]:5.1f}"
    for v in VALS: print(rows[v], flush=True)
    del ws; torch.cuda.empty_cache()
