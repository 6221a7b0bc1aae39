#!/usr/bin/env python
"""bench.py -- the headline benchmark of the Vec-LUT MUL_MAT path on B200 (driver contract in the task statement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--type I2_V_4] [--tokens 512]

Workload (BASELINE.json `metric`): ternary GEMM int-op/s (+ HBM GB/s) at K=4096 x N=14336 (Llama3-8B FFN up/gate
shape), I2_V_4 weights, synthetic random ternary weights and N(0,1) activations.
  * one *step* = one GGML_OP_MUL_MAT(vlut weight [K,N], f32 [K,T]) at T = 512 tokens (the pp512 regime):
    activation quantisation + Vec-LUT GEMM + fp32 epilogue;  value = 2*K*N*T / time  (zeros counted, like the
    reference's own test, tests/test-vlut-gemm.cpp:595-598), in Tint-op/s.
  * `sweep` = the same op at T in {1,2,4,...,2048}: per-T time, Tint-op/s, algorithmic HBM GB/s and both roofline
    fractions, measured in the same run with CUDA events.
  * inputs are resident in HBM for `value`; `e2e` is the same op through the C-ABI host-buffer entry
    (vlut_mul_mat_host: H2D of x from pinned memory, the op, D2H of the result, stream sync) -- what a
    ggml_backend_sched split costs.
  * L2: the timed loop rotates over enough distinct weight buffers (and activation/output buffers) that the working
    set exceeds the 126 MB L2 (`config.l2`).
  * N > 1 (torchrun): output-row tensor parallelism, weak scaling: every rank holds its own [K, 14336] row shard
    of a [K, 14336*N] matrix, quantises the (replicated) activations, computes its slice and all-gathers the f32
    output slices over NCCL/NVLink; value = 2*K*(14336*N)*T / max-over-ranks time.
  * --impl reference: the reference's OWN CPU kernels (oracle/_ref, compiled from /root/reference in place) on the
    box's host cores, same op, same metric.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

K_DIM, N_DIM = 4096, 14336
SWEEP_T = [1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048]
L2_BYTES = 126 * 1024 * 1024


def load_peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 100 ms while the timed regions run."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None
        self.marks = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ts, line in self.rows:
            f = [s.strip() for s in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # "under load" = the upper half of the samples (idle samples before/after the kernels sit at low clocks)
        sm_sorted = sorted(sm)
        load = sm_sorted[len(sm_sorted) // 2:]
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


def run_reference(args) -> None:
    """--impl reference: the reference's own CPU kernels on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    from oracle import oracle as O
    ref = O.RefLib()
    type_ = O.__dict__[args.type]
    T, K, N = args.tokens, K_DIM, N_DIM * args.gpus
    # all host cores, whatever OMP_NUM_THREADS says (torchrun exports OMP_NUM_THREADS=1)
    nth = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    rng = np.random.Generator(np.random.PCG64(1234))
    packed = rng.integers(0, 81 if type_ in (O.I2_V, O.I2_V_4, O.I2_V_8) else 243, size=O.row_size(type_, K) * N,
                          dtype=np.uint8)
    x = rng.standard_normal((T, K), dtype=np.float32)
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        ref.mul_mat(type_, packed, x, N, nth=nth)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt)
    total = sum(times)
    ops = 2.0 * K * N * T
    value = ops * len(times) / total / 1e12
    line = {
        "impl": "reference", "metric": "ternary_gemm_int_ops_per_s", "value": value, "unit": "Tint-op/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total / len(times) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8xternary->int32->f32",
        "data": "synthetic",
        "config": {"workload": f"vlut MUL_MAT {args.type} K={K} N={N} T={T} (reference CPU kernels, correct tile layout)",
                   "K": K, "N": N, "T": T, "type": args.type},
        "cpu_baseline": {"value": value, "unit": "Tint-op/s", "cores": nth, "kind": "reference",
                         "sample": f"full op, {len(times)} reps, isa build {ref.isa}"},
        "e2e": {"value": value, "unit": "Tint-op/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--type", default="I2_V_4", choices=["I2_V", "I2_V_4", "I2_V_8", "I1_V", "I1_V_2"])
    ap.add_argument("--tokens", type=int, default=512)
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--sweep-iters", type=int, default=60)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    if args.impl == "reference":
        run_reference(args)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    import vlut_b200  # noqa: F401
    from vlut_b200 import api

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the product path has no CPU fallback (use --impl reference for the "
                         "CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    peaks = load_peaks()
    int8_peak_tops = 2.0 * peaks["bf16_tflops"]  # dense int8 = 2x dense bf16 on sm_100a; no int8 figure is measured
    type_ = api.TYPE_BY_NAME[args.type]
    K, N, T = K_DIM, N_DIM, args.tokens
    nb_gguf = api.nbytes(type_, K, N)
    is_i2 = type_ in (api.I2_V, api.I2_V_4, api.I2_V_8)

    # a capturable (non-default) stream: launch-bound loops (two ~5 us kernels per op at small T against ~6 us of
    # Python/ctypes launch overhead per call) are captured once into CUDA graphs and replayed
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx = api.Context(dev, stream=stream)
    # the activations of every step are resident tensors written (and synchronised with) before the timed region
    ctx.set_option(api.Context.OPT_INPUTS_VISIBLE, 1)

    def graph_time_us(fn, iters, finish=None):
        """GPU time per call of fn(i), i in [0, iters): one CUDA graph of the whole loop, best of 3 replays, CUDA events
        on the launch stream.  Falls back to an eager loop if the capture fails."""
        for i in range(5):
            fn(i)
        torch.cuda.synchronize()
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                for i in range(iters):
                    fn(i)
                if finish:
                    finish()
            g.replay()
            torch.cuda.synchronize()
            best = None
            for _ in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                g.replay()
                e1.record()
                torch.cuda.synchronize()
                t = e0.elapsed_time(e1) / iters * 1e3
                best = t if best is None else min(best, t)
            del g
            return best, "cuda-graph"
        except Exception as e:  # noqa: BLE001
            print(f"[bench] graph capture failed ({e}); timing an eager loop", file=sys.stderr)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(iters):
                fn(i)
            if finish:
                finish()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / iters * 1e3, "eager"
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)

    # ---- resident inputs: enough distinct weight copies to exceed L2 -------------------------------------
    n_w = max(2, -(-int(1.25 * L2_BYTES) // nb_gguf))
    weights = []
    for i in range(n_w):
        # uniform bytes in [0, 81) (resp. [0, 243)) are iid uniform ternary digits: a valid GGUF I2/I1 tensor
        gg = torch.randint(0, 81 if is_i2 else 243, (nb_gguf,), dtype=torch.uint8, device=dev, generator=gen)
        weights.append(ctx.upload(type_, K, N, gg))
        del gg
    n_x = 4

    def make_x(t):
        g2 = torch.Generator(device=dev)
        g2.manual_seed(4321)  # activations are replicated across ranks
        return [torch.randn((t, K), dtype=torch.float32, device=dev, generator=g2) for _ in range(n_x)]

    xs = make_x(T)
    n_total = N * world
    gather_mode = "none"
    use_mc = False
    if world == 1:
        outs = [torch.empty((T, n_total), dtype=torch.float32, device=dev) for _ in range(n_x)]
    else:
        # fused all-gather: the GEMM epilogue stores into every rank's [T][N_total] symmetric-memory buffer (NVSwitch
        # multicast when available, peer pointers otherwise); NCCL all-gather + permute as the fallback / comparison
        outs, hdls = [], []
        try:
            if os.environ.get("VLUT_BENCH_GATHER", "fused") != "fused":
                raise RuntimeError("nccl gather requested")
            import torch.distributed._symmetric_memory as symm_mem
            for _ in range(n_x):
                buf = symm_mem.empty((T, n_total), dtype=torch.float32, device=dev)
                hdls.append(symm_mem.rendezvous(buf, dist.group.WORLD))
                outs.append(buf)
            # multicast (one multimem.st per element, replicated by the switch) is the default wherever it exists; plain
            # peer-pointer stores measured 185 us against 105 us per step on 2 x B200 (VLUT_BENCH_MULTICAST=0 selects them)
            use_mc = bool(int(getattr(hdls[0], "multicast_ptr", 0) or 0)) and os.environ.get("VLUT_BENCH_MULTICAST") != "0"
            gather_mode = "fused-multimem" if use_mc else "fused-p2p"
        except Exception as e:  # noqa: BLE001
            if rank == 0:
                print(f"[bench] symmetric memory unavailable ({e}); using the NCCL all-gather", file=sys.stderr)
            outs = [torch.empty((T, n_total), dtype=torch.float32, device=dev) for _ in range(n_x)]
            hdls = []
            gather_mode = "nccl"
        gather_buf = torch.empty((world * T, N), dtype=torch.float32, device=dev)
        local_out = [torch.empty((T, N), dtype=torch.float32, device=dev) for _ in range(n_x)]

    def step(i: int):
        w = weights[i % n_w]
        x = xs[i % n_x]
        if world == 1:
            ctx.mul_mat(w, x, out=outs[i % n_x])
        elif hdls:
            # the result buffers rotate: the cross-rank barrier that publishes step i-1 is issued after step i's GEMM
            # has been launched, so the NVLink transfer of one step overlaps the compute of the next
            h = hdls[i % n_x]
            ctx.mul_mat_scatter(w, x, peers=[int(p) for p in h.buffer_ptrs], multicast_ptr=int(getattr(h, "multicast_ptr", 0) or 0) if use_mc else 0,
                                ldo=n_total, col0=rank * N)
            if pending:
                pending.pop().barrier()
            pending.append(h)
        else:
            lo = local_out[i % n_x]
            ctx.mul_mat(w, x, out=lo)
            dist.all_gather_into_tensor(gather_buf, lo)
            outs[i % n_x].view(T, world, N).copy_(gather_buf.view(world, T, N).permute(1, 0, 2))

    pending = []  # symmetric-memory handles whose publishing barrier is still owed

    def barrier():
        while pending:
            pending.pop().barrier()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # quick in-run sanity: AUTO agrees with the scalar anchor kernel at the bench size (both on the GPU)
    torch.cuda.synchronize()
    chk_a = ctx.mul_mat(weights[0], xs[0]).clone()
    chk_b = ctx.mul_mat(weights[0], xs[0], algo=api.ALGO_SIMPLE)
    auto_equals_simple = bool(torch.equal(chk_a, chk_b))
    del chk_a, chk_b

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    for i in range(args.warmup):
        step(i)
    barrier()
    launches0 = ctx.launch_count
    timing_mode = "eager"
    if world == 1:
        # single GPU: the K steps are one CUDA graph (same kernels, same order, same rotating buffers)
        us, timing_mode = graph_time_us(step, args.steps)
        launches = (ctx.launch_count - launches0) * args.steps // (args.steps + 5) if timing_mode == "cuda-graph" else \
            (ctx.launch_count - launches0) * args.steps // (2 * args.steps + 5)
        total_ms = us * args.steps / 1e3
        per_step = [us / 1e3] * args.steps
    elif hdls and os.environ.get("VLUT_BENCH_GRAPH", "1") != "0":
        # fused gather: the K steps (GEMM with scatter epilogue + the symmetric-memory barrier that publishes each step)
        # are one CUDA graph per rank; the in-graph barriers keep the ranks in lockstep, time = max over ranks
        def flush():
            while pending:
                pending.pop().barrier()
        flush()
        dist.barrier()
        torch.cuda.synchronize()
        us, timing_mode = graph_time_us(step, args.steps, finish=flush)
        barrier()
        launches = (ctx.launch_count - launches0) * args.steps // (args.steps + 5) if timing_mode == "cuda-graph" else \
            (ctx.launch_count - launches0) * args.steps // (2 * args.steps + 5)
        total_ms = us * args.steps / 1e3
        per_step = [us / 1e3] * args.steps
    else:
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
        ev[0].record()
        for i in range(args.steps):
            step(i)
            ev[i + 1].record()
        barrier()
        launches = ctx.launch_count - launches0
        total_ms = ev[0].elapsed_time(ev[-1])
        per_step = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    if world > 1:
        tt = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms = float(tt.item())
    ms_per_step = total_ms / args.steps
    ops_step = 2.0 * K * n_total * T
    value = ops_step / (ms_per_step * 1e-3) / 1e12
    algo_used = api.ALGO_NAMES[ctx.last_algo]

    # ---- dominant kernel alone (GEMM+epilogue on pre-quantised activations), CUDA events on the launch stream ----
    def time_gemm(t, iters, x_list):
        ctx.quantize_act(type_, x_list[0])
        o = torch.empty((t, N), dtype=torch.float32, device=dev)
        us, _ = graph_time_us(lambda i: ctx.gemm(weights[i % n_w], t, out=o), iters)
        return us, api.ALGO_NAMES[ctx.last_algo]

    def time_op(t, iters, x_list):
        o = torch.empty((t, N), dtype=torch.float32, device=dev)
        us, _ = graph_time_us(lambda i: ctx.mul_mat(weights[i % n_w], x_list[i % n_x], out=o), iters)
        return us

    def alg_bytes(t):  # W (GGUF packed size) + A (f32 in) + O (f32 out)
        return nb_gguf + 4 * t * K + 4 * t * N

    gemm_us, gemm_algo = time_gemm(T, max(20, args.steps // 2), xs)
    gemm_tops = 2.0 * K * N * T / (gemm_us * 1e-6) / 1e12
    gemm_gbs = alg_bytes(T) / (gemm_us * 1e-6) / 1e9
    frac_tensor = gemm_tops / int8_peak_tops
    frac_hbm = gemm_gbs / peaks["hbm_gbs"]
    if frac_hbm >= frac_tensor:
        roofline = {"bound": "hbm", "achieved": gemm_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": frac_hbm}
    else:
        roofline = {"bound": "tensor", "achieved": gemm_tops, "peak": int8_peak_tops, "unit": "TOP/s",
                    "frac": frac_tensor}
    traffic = None
    try:
        tj = json.loads((ROOT / "profiles" / "r1_traffic.json").read_text())
        traffic = tj.get(f"{gemm_algo}_T{T}_{args.type}")
    except Exception:  # noqa: BLE001
        pass
    roofline.update({"traffic": traffic, "kernel": f"vlut gemm ({gemm_algo}) T={T}", "kernel_us": gemm_us,
                     "peak_source": f"{peaks['source']} (MEASURED_PEAKS.json); int8 tensor peak taken as 2 x measured "
                                    f"bf16 burst (no int8 figure is measured)",
                     "algorithmic_bytes": alg_bytes(T), "algorithmic_ops": 2 * K * N * T})

    sweep = []
    if not args.no_sweep and world == 1:
        for t in SWEEP_T:
            xl = make_x(t)
            it = args.sweep_iters if t <= 256 else max(10, args.sweep_iters // 4)
            g_us, g_algo = time_gemm(t, it, xl)
            op_us = time_op(t, it, xl)
            tops = 2.0 * K * N * t / (g_us * 1e-6) / 1e12
            gbs = alg_bytes(t) / (g_us * 1e-6) / 1e9
            sweep.append({"T": t, "algo": g_algo, "gemm_us": round(g_us, 3), "op_us": round(op_us, 3),
                          "tint_ops": round(tops, 3), "alg_gbs": round(gbs, 1),
                          "frac_hbm": round(gbs / peaks["hbm_gbs"], 4), "frac_int8": round(tops / int8_peak_tops, 4),
                          "tok_s": round(t / (op_us * 1e-6), 1)})
            del xl

    # ---- e2e through the C-ABI host-buffer entry ----------------------------------------------------------
    xh = [x.cpu().pin_memory() for x in xs[:2]]
    oh = [torch.empty((T, N), dtype=torch.float32).pin_memory() for _ in range(2)]
    e2e_steps = max(5, min(args.steps, 50))
    for i in range(3):
        ctx.mul_mat_host(weights[i % n_w], xh[i % 2], oh[i % 2])
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        ctx.mul_mat_host(weights[i % n_w], xh[i % 2], oh[i % 2])
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        tt = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_s = float(tt.item())
    e2e = {"value": 2.0 * K * n_total * T / e2e_s / 1e12, "unit": "Tint-op/s", "h2d_bytes_per_step": 4 * T * K,
           "d2h_bytes_per_step": 4 * T * N, "ms_per_step": e2e_s * 1e3,
           "note": "vlut_mul_mat_host: pinned H2D of x + quantise + GEMM + D2H of out + stream sync, per rank"}

    clocks = sampler.stop() if rank == 0 else None

    # ---- CPU baseline on the host cores (rank 0, N=1 only) --------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import oracle as O
        try:
            ref = O.RefLib()
            kind = "reference"
            nth = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
            pk = weights[0]  # regenerate the same distribution on the host: timing is data independent
            rng = np.random.Generator(np.random.PCG64(1234))
            packed = rng.integers(0, 81 if is_i2 else 243, size=nb_gguf, dtype=np.uint8)
            xc = xs[0].cpu().numpy()
            ref.mul_mat(type_, packed, xc, N, nth=nth)
            reps, t_acc = 0, 0.0
            while t_acc < 10.0 and reps < 20:
                t0 = time.perf_counter()
                ref.mul_mat(type_, packed, xc, N, nth=nth)
                t_acc += time.perf_counter() - t0
                reps += 1
            cpu = {"value": 2.0 * K * N * T * reps / t_acc / 1e12, "unit": "Tint-op/s", "cores": nth, "kind": kind,
                   "sample": f"full op K={K} N={N} T={T}, {reps} reps, {t_acc / reps * 1e3:.1f} ms each, "
                             f"build {ref.isa}", "ms_per_op": t_acc / reps * 1e3}
        except Exception as e:  # noqa: BLE001
            cpu = {"value": None, "unit": "Tint-op/s", "cores": 0, "kind": "reference", "sample": f"failed: {e}"}

    if rank == 0:
        ws_mb = (n_w * api.repacked_nbytes(type_, K, N) + n_x * 4 * T * (K + n_total)) / 1e6
        line = {
            "metric": "ternary_gemm_int_ops_per_s", "value": value, "unit": "Tint-op/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int8xternary->int32->f32", "data": "synthetic",
            "config": {"workload": f"vlut MUL_MAT {args.type} K={K} N={n_total} T={T} "
                                   f"(quantise + Vec-LUT GEMM + fp32 epilogue"
                                   f"{', row-sharded + all-gather of output slices' if world > 1 else ''})",
                       "K": K, "N": n_total, "T": T, "type": args.type, "algo": algo_used,
                       "l2": f"rotating {n_w} weight buffers + {n_x} activation/output buffers, working set "
                             f"{ws_mb:.0f} MB > 126 MB L2",
                       "parallelism": f"row-shard x{world}, gather={gather_mode}" if world > 1 else "single GPU"},
            "median_ms_per_step": statistics.median(per_step),
            "timing": f"{timing_mode}: K steps timed with CUDA events on the launch stream",
            "gpu_launches": int(launches),
            "roofline": roofline,
            "cpu_baseline": cpu,
            "e2e": e2e,
            "clocks": clocks,
            "sweep": sweep,
            "checks": {"auto_equals_simple_kernel": auto_equals_simple},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
