"""sweep_small.py -- small-T experiments: GEMM-only and quantise+GEMM time per (shape, T, algo, env knobs).

Weights rotate over more than L2 (126 MB) worth of distinct buffers, CUDA events on the launch stream, 5 warm-up + ITERS timed
launches per point.  Knobs are environment variables the launchers read at every call (VLUT_TCS_SPLIT, VLUT_TCS_DEPTH, ...).

usage: python scripts/sweep_small.py [quick|full|quant]
"""
import os
import sys

sys.path.insert(0, ".")
import torch

import vlut_b200  # noqa: F401
from vlut_b200 import api

ITERS = 100
L2 = 126 * 1024 * 1024
mode = sys.argv[1] if len(sys.argv) > 1 else "quick"
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)   # a capturable (non-default) stream: the timed loops are replayed as CUDA graphs
torch.cuda.set_stream(stream)
ctx = api.Context(dev, stream=stream)
gen = torch.Generator(device=dev)
gen.manual_seed(7)


def weights_for(type_, K, N):
    nb = api.nbytes(type_, K, N)
    n_w = max(2, -(-int(1.25 * L2) // nb))
    n_w = min(n_w, 40)
    hi = 81 if type_ in (api.I2_V, api.I2_V_4, api.I2_V_8) else 243
    ws = []
    for _ in range(n_w):
        gg = torch.randint(0, hi, (nb,), dtype=torch.uint8, device=dev, generator=gen)
        ws.append(ctx.upload(type_, K, N, gg))
    return ws, nb


def timed(fn, iters=ITERS):
    """us per call, GPU time: the loop is captured once into a CUDA graph (launch overhead of the Python/ctypes caller,
    ~6 us per call, would otherwise hide every kernel shorter than that) and replayed."""
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    if os.environ.get("SWEEP_EAGER"):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters * 1e3
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        for i in range(iters):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / iters * 1e3)
    del g
    return best


def setenv(env):
    for k in ("VLUT_TCS_SPLIT", "VLUT_TCS_DEPTH", "VLUT_TCS_NAS", "VLUT_TCS_PAIR", "VLUT_PDL", "VLUT_QUANT_2PASS"):
        os.environ.pop(k, None)
    for k, v in env.items():
        os.environ[k] = str(v)


def sweep(type_, K, N, Ts, configs):
    ws, nb = weights_for(type_, K, N)
    n_w = len(ws)
    for T in Ts:
        xs = [torch.randn((T, K), dtype=torch.float32, device=dev, generator=gen) for _ in range(4)]
        o = torch.empty((T, N), dtype=torch.float32, device=dev)
        ref = None
        for name, algo, env in configs:
            setenv(env)
            try:
                ctx.quantize_act(type_, xs[0])
                g_us = timed(lambda i: ctx.gemm(ws[i % n_w], T, out=o, algo=algo))
                op_us = timed(lambda i: ctx.mul_mat(ws[i % n_w], xs[i % 4], out=o, algo=algo))
                res = ctx.mul_mat(ws[0], xs[0], algo=algo).clone()
                if ref is None:
                    ref = res
                same = bool(torch.equal(res, ref))
            except Exception as e:  # noqa: BLE001
                print(f"{api.TYPE_NAMES.get(type_, type_) if hasattr(api, 'TYPE_NAMES') else type_} K={K} N={N} T={T} {name}: ERROR {e}", flush=True)
                continue
            bytes_ = nb + 4 * T * K + 4 * T * N
            print(f"type={type_} K={K} N={N} T={T:3d} {name:28s} gemm {g_us:6.2f} us ({bytes_ / g_us / 1e3:6.0f} GB/s)  "
                  f"op {op_us:6.2f} us  same={same}", flush=True)
    setenv({})
    del ws
    torch.cuda.empty_cache()


TC, TCS = api.ALGO_TC, api.ALGO_TCS
if mode == "quant":
    # quantise kernel alone: old two-pass vs register-resident, with and without PDL
    for K in (4096, 14336):
        for T in (1, 16, 64, 512, 2048):
            x = torch.randn((T, K), dtype=torch.float32, device=dev, generator=gen)
            for name, env in (("two-pass", {"VLUT_QUANT_2PASS": 1}), ("register", {}), ("register,nopdl", {"VLUT_PDL": 0})):
                setenv(env)
                us = timed(lambda i: ctx.quantize_act(api.I2_V_4, x))
                print(f"quantise K={K} T={T:4d} {name:16s} {us:6.2f} us  ({(4 * T * K + 2 * T * K) / us / 1e3:6.0f} GB/s)", flush=True)
    sys.exit(0)

base = [("tc (old)", TC, {}), ("tc (old), 2pass quant", TC, {"VLUT_QUANT_2PASS": 1})]
tcs_cfgs = [("tcs auto", TCS, {})]
for s in (1, 2, 4, 8):
    tcs_cfgs.append((f"tcs split={s}", TCS, {"VLUT_TCS_SPLIT": s}))
tcs_cfgs.append(("tcs split=1 depth=8", TCS, {"VLUT_TCS_SPLIT": 1, "VLUT_TCS_DEPTH": 8}))
tcs_cfgs.append(("tcs split=2 depth=8", TCS, {"VLUT_TCS_SPLIT": 2, "VLUT_TCS_DEPTH": 8}))
tcs_cfgs.append(("tcs split=1 nopdl", TCS, {"VLUT_TCS_SPLIT": 1, "VLUT_PDL": 0}))

if mode == "quick":
    sweep(api.I2_V_4, 4096, 14336, [1, 16, 32, 64], base + tcs_cfgs)
    sweep(api.I2_V_4, 4096, 4096, [16], base + tcs_cfgs)
else:
    sweep(api.I2_V_4, 4096, 14336, [1, 4, 16, 32, 64], base + tcs_cfgs)
    sweep(api.I2_V_4, 4096, 4096, [1, 16, 64], base + tcs_cfgs)
    sweep(api.I2_V_4, 14336, 4096, [1, 16, 64], base + tcs_cfgs)
    sweep(api.I2_V_4, 4096, 1024, [16], base + tcs_cfgs)
    sweep(api.I1_V_2, 3200, 8640, [1, 16, 64], base + tcs_cfgs)
    sweep(api.I1_V_2, 8640, 3200, [16], base + tcs_cfgs)
