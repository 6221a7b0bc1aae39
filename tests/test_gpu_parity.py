"""GPU suite: the CUDA path, called through the C ABI (libvlut_b200.so via ctypes), against the oracle.

Bit-exact everywhere (the op is integer arithmetic followed by one fp32 multiply): fp32 outputs are compared as
uint32 bit patterns -- a stricter bar than north_star's 1e-6 relative tolerance.
"""
import glob
import hashlib
import json
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

GOLDEN = Path(__file__).resolve().parent / "golden"
FIXTURES = sorted(glob.glob(str(GOLDEN / "*.npz")))


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def api():
    import vlut_b200  # noqa: F401
    from vlut_b200 import api as a
    return a


def algos_for(T):
    a = api()
    return ([a.ALGO_AUTO, a.ALGO_SIMPLE] + ([a.ALGO_LUT] if T <= 16 else []) +
            ([a.ALGO_GEMV] if T <= 8 else []) +   # CUDA-core split-K GEMV with fused quantisation
            ([a.ALGO_TCP] if T > 32 else []) +    # CTA-pair kernel: token tiles of 64 / 128 / 256
            ([a.ALGO_TCS] if T <= 64 else []))    # small-T streaming kernel


def run(ctx, type_, packed, x, N, algo, w_scale=1.0):
    a = api()
    T, K = x.shape
    w = ctx.upload(type_, K, N, packed)
    xd = torch.from_numpy(x).cuda()
    out = ctx.mul_mat(w, xd, w_scale=w_scale, algo=algo)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_native_library_is_loaded(ctx):
    import vlut_b200
    maps = open("/proc/self/maps").read()
    assert str(vlut_b200.LIB_PATH) in maps
    assert api().device_count() >= 1


@pytest.mark.parametrize("type_,K,N", [(O.I2_V, 1024, 130), (O.I2_V_4, 4096, 256), (O.I2_V_8, 2048, 64),
                                       (O.I1_V, 1300, 77), (O.I1_V_2, 3200, 128), (O.I1_V_2, 4096, 40),
                                       (O.I1_V, 1048, 16)])
def test_repack_round_trip(ctx, orc, type_, K, N):
    """GGUF bytes -> T2 (device) -> ternary == the weights that were packed (all five packings, ragged N, I1 tail)."""
    w = O.make_ternary(N, K, seed=11)
    packed = orc.pack(type_, w)
    W = ctx.upload(type_, K, N, packed)
    assert np.array_equal(ctx.unpack(W), w.astype(np.int8))
    # device-resident source takes the same path
    W2 = ctx.upload(type_, K, N, torch.from_numpy(packed).cuda())
    torch.cuda.synchronize()
    assert torch.equal(W.t2, W2.t2)


# VLUT_QUANT_KERNEL / VLUT_QUANT_THREADS: register-resident rows with 128 / 256 / 512 threads per token, or the two-pass kernel
QUANT_KERNELS = {"register-kernel": (0, 0), "register-128": (0, 128), "register-256": (0, 256), "register-512": (0, 512),
                 "two-pass-kernel": (2, 0)}


@pytest.mark.parametrize("kernel", list(QUANT_KERNELS.values()), ids=list(QUANT_KERNELS))
@pytest.mark.parametrize("kind", ["normal", "uniform", "outlier", "negative", "zeros", "ones"])
@pytest.mark.parametrize("T,K", [(1, 512), (7, 4096), (33, 3200), (70, 1300), (5, 8640), (18, 14336), (3, 16400), (600, 2048),
                                 (2, 32768)])
def test_quantize_act_bit_exact(ctx, orc, knobs, kernel, kind, T, K):
    if kernel != (0, 0) and kind not in ("normal", "outlier", "negative"):
        pytest.skip("three kinds are enough for the non-default launches")
    knobs(VLUT_QUANT_KERNEL=kernel[0], VLUT_QUANT_THREADS=kernel[1])
    x = O.make_acts(T, K, seed=T + K, kind=kind)
    q_ref, s_ref = orc.quantize(x)
    ctx.quantize_act(O.I2_V_4 if K % 16 == 0 else O.I1_V, torch.from_numpy(x).cuda())
    q, s = ctx.get_act(K, T)
    assert np.array_equal(q, q_ref)
    assert np.array_equal(bits(s), bits(s_ref))


@pytest.mark.parametrize("two_pass", list(QUANT_KERNELS.values()), ids=list(QUANT_KERNELS))
def test_quantize_ties_and_saturation(ctx, orc, knobs, two_pass):
    """round() is half AWAY from zero and the clamp is asymmetric: exact .5 products (row max 127 -> scale 1), values far
    outside the int8 range, and a row whose signed max is negative (everything saturates to -128)."""
    knobs(VLUT_QUANT_KERNEL=two_pass[0], VLUT_QUANT_THREADS=two_pass[1])
    K = 1024
    x = np.zeros((3, K), dtype=np.float32)
    ties = np.array([0.5, -0.5, 1.5, -1.5, 2.5, -2.5, 126.5, -126.5, -127.5, -128.5, -129.0, -1000.0, 127.0, 0.49999997,
                     -0.49999997, 63.5], dtype=np.float32)
    x[0, :ties.size] = ties
    x[0, 500] = 127.0
    x[1] = np.linspace(-300, 127, K, dtype=np.float32)
    x[2] = -np.abs(O.make_acts(1, K, seed=5)[0]) - 0.25
    q_ref, s_ref = orc.quantize(x)
    ctx.quantize_act(O.I2_V_4, torch.from_numpy(x).cuda())
    q, s = ctx.get_act(K, 3)
    assert np.array_equal(q, q_ref)
    assert np.array_equal(bits(s), bits(s_ref))
    assert list(q[0, :6]) == [1, -1, 2, -2, 3, -3]


def test_quantize_strided_rows(ctx, orc):
    # ggml's src1 may have nb11 > 4*K
    T, K, ld = 9, 1024, 1100
    buf = np.zeros((T, ld), dtype=np.float32)
    buf[:, :K] = O.make_acts(T, K, seed=1)
    xd = torch.from_numpy(buf).cuda()[:, :K]
    ctx.quantize_act(O.I2_V_4, xd)
    q, s = ctx.get_act(K, T)
    q_ref, s_ref = orc.quantize(buf[:, :K])
    assert np.array_equal(q, q_ref) and np.array_equal(bits(s), bits(s_ref))
    # rows that are not 16-byte aligned take the two-pass kernel
    xd2 = torch.from_numpy(np.ascontiguousarray(np.pad(buf, ((0, 0), (1, 0)))) ).cuda()[:, 1:K + 1]
    ctx.quantize_act(O.I2_V_4, xd2)
    q2, s2 = ctx.get_act(K, T)
    assert np.array_equal(q2, q_ref) and np.array_equal(bits(s2), bits(s_ref))


@pytest.mark.parametrize("path", FIXTURES, ids=lambda p: Path(p).stem)
def test_golden_fixtures(ctx, path):
    """Committed outputs of the reference's own kernels (tests/golden, oracle/make_golden.py)."""
    z = np.load(path)
    type_, K, N, T = int(z["type"]), int(z["K"]), int(z["N"]), int(z["T"])
    for algo in algos_for(T):
        out = run(ctx, type_, z["packed"], z["x"], N, algo)
        assert np.array_equal(bits(out), bits(z["out"])), f"algo={algo}"
    q, s = ctx.get_act(K, T)
    assert np.array_equal(q, z["q"]) and np.array_equal(bits(s), bits(z["scales"]))


SHAPES = [
    # type, K, N  -- model shapes scaled down in N, plus ragged / tail cases
    (O.I2_V_4, 4096, 384), (O.I2_V_4, 2048, 200), (O.I2_V, 1040, 129), (O.I2_V_8, 3072, 256),
    (O.I1_V_2, 3200, 320), (O.I1_V_2, 8640, 136), (O.I1_V, 1300, 64), (O.I1_V_2, 4096, 128), (O.I1_V, 3072, 72),
]


@pytest.mark.parametrize("T", [1, 2, 7, 16, 17, 32, 33, 64, 100, 256])
@pytest.mark.parametrize("type_,K,N", SHAPES)
def test_mul_mat_bit_exact(ctx, orc, type_, K, N, T):
    w = O.make_ternary(N, K, seed=K + N)
    x = O.make_acts(T, K, seed=T, kind="outlier" if T % 2 else "normal")
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    for algo in algos_for(T):
        out = run(ctx, type_, packed, x, N, algo)
        bad = bits(out) != bits(ref)
        assert not bad.any(), f"algo={algo}: {bad.sum()} of {bad.size} differ, first at {np.argwhere(bad)[0]}"


@pytest.mark.parametrize("type_,K", [(O.I2_V_4, 4096), (O.I1_V_2, 3200), (O.I2_V_4, 1536)])
@pytest.mark.parametrize("T", [3, 40, 200])
def test_int16_wrap(ctx, orc, type_, K, T):
    """Adversarial: all-(+1) weights x saturated activations overflow int16 in every K-group."""
    N = 144
    w = np.ones((N, K), dtype=np.float32)
    w[::3] = -1.0
    x = O.make_acts(T, K, kind="ones")
    x[1::2] *= -1.0  # negative rows saturate to -128
    x[:, ::5] = 0.5
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    for algo in algos_for(T):
        out = run(ctx, type_, packed, x, N, algo)
        assert np.array_equal(bits(out), bits(ref)), f"algo={algo}"


def test_full_size_known_answers(ctx):
    """BASELINE.json configs at full size: sha256 of the fp32 result == the reference's own (kat_full.json)."""
    kats = json.loads((GOLDEN / "kat_full.json").read_text())
    for name, k in kats.items():
        w = O.make_ternary(k["N"], k["K"], seed=k["w_seed"])
        x = O.make_acts(k["T"], k["K"], seed=k["x_seed"], kind=k["x_kind"])
        if hashlib.sha256(x.tobytes()).hexdigest() != k["sha256_x"]:
            pytest.skip("numpy RNG stream differs from the one the fixtures were made with")
        packed = O.Oracle().pack(k["type"], w)
        assert hashlib.sha256(packed.tobytes()).hexdigest() == k["sha256_packed"]
        for algo in [api().ALGO_AUTO] + ([api().ALGO_TCS] if k["T"] <= 64 else [api().ALGO_TCP]):
            out = run(ctx, k["type"], packed, x, k["N"], algo)
            assert hashlib.sha256(out.tobytes()).hexdigest() == k["sha256_out"], f"{name} algo={algo}"


@pytest.mark.parametrize("T", [4, 48, 512])
def test_properties_at_full_size(ctx, T):
    """Size-independent properties at the headline shape (4096 -> 14336): determinism, agreement between
    algorithms, w_scale folding, row-shard concatenation == full result."""
    a = api()
    K, N = 4096, 14336
    rng = np.random.Generator(np.random.PCG64(T))
    w = (rng.integers(0, 3, size=(N, K), dtype=np.int8) - 1).astype(np.float32)
    packed = O.Oracle().pack(O.I2_V_4, w)
    x = rng.standard_normal((T, K), dtype=np.float32)
    W = ctx.upload(O.I2_V_4, K, N, packed)
    xd = torch.from_numpy(x).cuda()
    base = ctx.mul_mat(W, xd, algo=a.ALGO_AUTO).clone()
    again = ctx.mul_mat(W, xd, algo=a.ALGO_AUTO)
    assert torch.equal(base, again)  # deterministic (stream-K partials are integers)
    simple = ctx.mul_mat(W, xd, algo=a.ALGO_SIMPLE)
    assert torch.equal(base, simple)
    # exact integer check against torch on the device: out == (q @ w^T) * scale  (no wrap for these inputs)
    q, s = ctx.get_act(K, T)
    acc = (torch.from_numpy(q).cuda().double() @ torch.from_numpy(w).cuda().double().T)
    expect = (acc.float() * torch.from_numpy(s).cuda()[:, None])
    assert torch.equal(base, expect)
    # weight scale folded into the epilogue == a following MUL by the scalar (bitnet.cpp:33-34)
    scaled = ctx.mul_mat(W, xd, w_scale=0.0371, algo=a.ALGO_AUTO)
    assert torch.equal(scaled, base * torch.tensor(0.0371, dtype=torch.float32, device="cuda"))
    # output-row sharding: 4 shards concatenated == full
    parts = []
    for g in range(4):
        r0, rows = g * N // 4, N // 4
        Wg = ctx.upload(O.I2_V_4, K, rows, a.shard_rows(O.I2_V_4, K, N, r0, rows, packed))
        parts.append(ctx.mul_mat(Wg, xd, algo=a.ALGO_AUTO).clone())
    assert torch.equal(torch.cat(parts, dim=1), base)


def test_mul_mat_host_buffers(ctx, orc):
    a = api()
    type_, K, N, T = O.I2_V_4, 2048, 512, 24
    w = O.make_ternary(N, K, seed=1)
    x = O.make_acts(T, K, seed=2)
    packed = orc.pack(type_, w)
    W = ctx.upload(type_, K, N, packed)
    xh = torch.from_numpy(x).pin_memory()
    oh = torch.empty((T, N), dtype=torch.float32).pin_memory()
    ctx.mul_mat_host(W, xh, oh)
    assert np.array_equal(bits(oh.numpy()), bits(orc.mul_mat(type_, packed, x, N)))


@pytest.mark.parametrize("T", [129, 300, 512])
def test_mul_mat_host_pipeline_chunks(ctx, orc, T):
    """T > 128 takes the chunked H2D / compute / D2H pipeline of vlut_mul_mat_host (ragged last chunk, strided host
    rows, pageable and pinned memory, back-to-back calls reusing the staging buffers)."""
    type_, K, N = O.I2_V_4, 2048, 384
    w = O.make_ternary(N, K, seed=3)
    packed = orc.pack(type_, w)
    W = ctx.upload(type_, K, N, packed)
    for rep, pinned in enumerate([True, False, True]):
        x = O.make_acts(T, K, seed=10 + rep, kind="outlier" if rep == 1 else "normal")
        xh = torch.zeros((T, K + 16), dtype=torch.float32)
        oh = torch.full((T, N + 8), -3.0, dtype=torch.float32)
        if pinned:
            xh, oh = xh.pin_memory(), oh.pin_memory()
        xh[:, :K] = torch.from_numpy(x)
        ctx.mul_mat_host(W, xh[:, :K], oh[:, :N])
        assert np.array_equal(bits(oh[:, :N].numpy()), bits(orc.mul_mat(type_, packed, x, N)))
        assert (oh[:, N:] == -3.0).all()


@pytest.mark.parametrize("T", [8, 64, 512, 700])
def test_inputs_visible_option_and_chained_matmuls(orc, T):
    """VLUT_OPT_INPUTS_VISIBLE lets the quantise kernel read x ahead of its dependency wait while the previous GEMM of
    the context still runs -- except when that GEMM's output IS the next x (chained matmuls), which the library must
    detect.  Back-to-back ops without host synchronisation, K == N so that outputs can feed the next op."""
    a = api()
    c = a.Context(0)
    c.set_option(a.Context.OPT_INPUTS_VISIBLE, 1)
    type_, K = O.I2_V_4, 1024
    ws = [O.make_ternary(K, K, seed=s) for s in (1, 2)]
    packed = [orc.pack(type_, w) for w in ws]
    W = [c.upload(type_, K, K, p) for p in packed]
    x0 = O.make_acts(T, K, seed=3)
    xd = torch.from_numpy(x0).cuda()
    y1 = torch.empty((T, K), device="cuda")
    y2 = torch.empty((T, K), device="cuda")
    outs = []
    for rep in range(3):  # independent x (early allowed), then chained: y1 -> y2 (early must be refused)
        c.mul_mat(W[0], xd, out=y1)
        c.mul_mat(W[1], y1, out=y2)
        c.mul_mat(W[0], xd, out=y1)  # overwrites y1 while nothing reads it any more
        outs.append((y1.clone(), y2.clone()))
    torch.cuda.synchronize()
    r1 = orc.mul_mat(type_, packed[0], x0, K)
    r2 = orc.mul_mat(type_, packed[1], r1, K)
    for y1c, y2c in outs:
        assert np.array_equal(bits(y1c.cpu().numpy()), bits(r1))
        assert np.array_equal(bits(y2c.cpu().numpy()), bits(r2))
    c.close()


def test_scale_rows_matches_fp32_multiply(ctx):
    """vlut_scale_rows = ggml's MUL by a one-element f32 tensor: one fp32 rounding per element, strided rows, in place."""
    rng = np.random.Generator(np.random.PCG64(5))
    for T, N, s in [(1, 7, 0.37), (33, 1000, -1.9), (300, 4097, 3.0e-3)]:
        x = (rng.standard_normal((T, N + 5)) * 100).astype(np.float32)
        xd = torch.from_numpy(x).cuda()
        out = torch.full((T, N + 3), -7.0, device="cuda")
        ctx.scale_rows(xd[:, :N], s, out=out[:, :N])
        torch.cuda.synchronize()
        expect = x[:, :N] * np.float32(s)
        assert np.array_equal(bits(out[:, :N].cpu().numpy()), bits(expect))
        assert (out[:, N:] == -7.0).all()
        ctx.scale_rows(xd[:, :N], s, out=xd[:, :N])  # in place
        torch.cuda.synchronize()
        assert np.array_equal(bits(xd[:, :N].cpu().numpy()), bits(expect))
        assert np.array_equal(bits(xd[:, N:].cpu().numpy()), bits(x[:, N:]))


def test_strided_output_and_errors(ctx, orc):
    a = api()
    type_, K, N, T = O.I1_V_2, 1280, 100, 5
    w = O.make_ternary(N, K, seed=1)
    x = O.make_acts(T, K, seed=2)
    packed = orc.pack(type_, w)
    W = ctx.upload(type_, K, N, packed)
    big = torch.full((T, N + 28), -7.0, device="cuda")
    ctx.mul_mat(W, torch.from_numpy(x).cuda(), out=big[:, :N])
    torch.cuda.synchronize()
    assert np.array_equal(bits(big[:, :N].cpu().numpy()), bits(orc.mul_mat(type_, packed, x, N)))
    assert (big[:, N:] == -7.0).all()  # nothing written past N
    from vlut_b200._lib import VlutError
    with pytest.raises(VlutError):
        ctx.mul_mat(W, torch.from_numpy(x).cuda(), algo=99)
    with pytest.raises(ValueError):
        ctx.upload(O.I2_V_4, 1000, 8, np.zeros(10, dtype=np.uint8))  # K % 16 != 0


@pytest.mark.parametrize("nt", [0, 256], ids=["tile-auto", "tile-256"])
@pytest.mark.parametrize("fast", [1, 0], ids=["fastpath-allowed", "per-group-drain"])
@pytest.mark.parametrize("grid,stages", [(4, 0), (6, 2), (20, 3), (148, 0)])
@pytest.mark.parametrize("T", [33, 64, 130, 300, 520])
def test_streamk_split_tiles_and_ring_wrap(ctx, orc, knobs, grid, stages, T, fast, nt):
    """Force few persistent CTA pairs / few pipeline stages so that every pair walks several (tile, K-group) units:
    exercises the shared-memory ring wrap-around, tiles completed by a single pair (direct store), tiles split
    across pairs (int32 partials + last-arriver epilogue), the self-cleaning workspace and both drain policies (single
    drain when no int16 wrap is possible / per K-group)."""
    a = api()
    if nt == 256 and T <= 64:
        pytest.skip("the 256-token tile is for T > 64")
    knobs(VLUT_TC_GRID=grid, VLUT_TC_STAGES=stages, VLUT_TC_FAST=fast, VLUT_TC_NT=nt)
    for type_, K, N in [(O.I2_V_4, 4096, 640), (O.I1_V_2, 3200, 520)]:
        w = O.make_ternary(N, K, seed=grid)
        x = O.make_acts(T, K, seed=T, kind="uniform" if T % 2 else "normal")  # uniform: sum|q| often > 32767
        packed = orc.pack(type_, w)
        ref = orc.mul_mat(type_, packed, x, N)
        for rep in range(2):  # second run sees the workspace the first one left behind
            out = run(ctx, type_, packed, x, N, a.ALGO_TCP)
            assert np.array_equal(bits(out), bits(ref)), f"rep={rep}"


@pytest.mark.parametrize("nt", [0, 256], ids=["tile-auto", "tile-256"])
@pytest.mark.parametrize("fast", [1, 0], ids=["fastpath-allowed", "per-group-drain"])
@pytest.mark.parametrize("grid", [4, 10, 14, 22, 148])
@pytest.mark.parametrize("T", [65, 300, 600])
def test_tcp_hybrid_whole_plus_split_tiles(ctx, orc, knobs, grid, T, fast, nt):
    """VLUT_TC_WHOLE=2: every CTA pair walks its share of the split units first (tiles shared with other pairs; pairs
    with no split unit, ranges crossing a tile boundary, last-arriver fix-up skipping empty pairs) and then its whole
    tiles.  Forced grids make floor(tiles / pairs) and the remainder take many values."""
    a = api()
    knobs(VLUT_TC_GRID=grid, VLUT_TC_WHOLE=2, VLUT_TC_FAST=fast, VLUT_TC_NT=nt)
    for type_, K, N in [(O.I2_V_4, 2048, 1024), (O.I1_V_2, 3200, 900), (O.I2_V_4, 4096, 1800)]:
        w = O.make_ternary(N, K, seed=grid)
        x = O.make_acts(T, K, seed=T, kind="uniform" if grid % 4 else "normal")
        packed = orc.pack(type_, w)
        ref = orc.mul_mat(type_, packed, x, N)
        for rep in range(2):
            out = run(ctx, type_, packed, x, N, a.ALGO_TCP)
            assert np.array_equal(bits(out), bits(ref)), f"{O.TYPE_NAMES[type_]} K={K} N={N} rep={rep}"


@pytest.mark.parametrize("coop", [(1, 400000), (1, 1), (1, 3000), (0, 0)],
                         ids=["shared-fixup", "shared-fixup-everybody-gives-up", "shared-fixup-some-give-up", "last-arriver-fixup"])
@pytest.mark.parametrize("fast", [1, 0], ids=["fastpath-allowed", "per-group-drain"])
@pytest.mark.parametrize("grid", [12, 22, 148])
@pytest.mark.parametrize("T", [130, 300, 600])
def test_tcp_hybrid_shared_fixup(ctx, orc, knobs, grid, T, fast, coop):
    """Hybrid split with at most one split unit per pair (fewer split units than pairs): VLUT_TC_COOP=1 lets every contributor
    of a split tile wait for the others and finish its own 16-column chunks; 0 is the last-arriver fix-up.  The wait is bounded
    (VLUT_TC_COOP_TIMEOUT clk): with 1 clk every waiter gives up at once and the last arriver finishes the tile alone, with
    3000 clk some do and some do not.  More contributors
    than chunks (I1 flat walk: 25 blocks, 8 chunks), fewer (K = 1024: 2 units), ragged N and T, exact and fast path, twice in
    a row (the counters and accumulation tiles must come back clean)."""
    a = api()
    knobs(VLUT_TC_GRID=grid, VLUT_TC_WHOLE=2, VLUT_TC_FAST=fast, VLUT_TC_COOP=coop[0], VLUT_TC_COOP_TIMEOUT=coop[1])
    for type_, K, N in [(O.I2_V_4, 1024, 700), (O.I1_V_2, 3200, 520), (O.I2_V_4, 4096, 1300), (O.I1_V, 1300, 260)]:
        w = O.make_ternary(N, K, seed=grid + coop[0])
        x = O.make_acts(T, K, seed=T, kind="uniform" if (grid + T) % 4 == 0 else "normal")
        packed = orc.pack(type_, w)
        ref = orc.mul_mat(type_, packed, x, N)
        for rep in range(2):
            out = run(ctx, type_, packed, x, N, a.ALGO_TCP)
            assert np.array_equal(bits(out), bits(ref)), f"{O.TYPE_NAMES[type_]} K={K} N={N} rep={rep}"


@pytest.mark.parametrize("fast", [1, 0], ids=["fastpath-allowed", "per-group-drain"])
@pytest.mark.parametrize("split", [1, 2, 3, 5, 8])
@pytest.mark.parametrize("T", [1, 9, 16, 31, 64])
def test_tcs_cluster_split_k(ctx, orc, knobs, split, T, fast):
    """VLUT_ALGO_TCS with the K-groups of every strip forced onto `split` CTAs of a thread-block cluster (capped at the
    number of K-groups): distributed-shared-memory reduction, uneven group and token slices, activation-ring wrap-around
    (VLUT_TCS_NAS=2), both drain policies, I2 (512-k groups) and I1 (640-k groups, odd sub-block count, 4-weight tail)."""
    a = api()
    knobs(VLUT_TCS_SPLIT=split, VLUT_TCS_NAS=2 if split % 2 else 0, VLUT_TC_FAST=fast)
    for type_, K, N in [(O.I2_V_4, 4096, 640), (O.I1_V_2, 3200, 520), (O.I2_V, 1040, 129), (O.I1_V, 3072, 72),
                        (O.I2_V_4, 14336, 256)]:
        w = O.make_ternary(N, K, seed=split)
        x = O.make_acts(T, K, seed=T, kind="uniform" if T % 2 else "normal")  # uniform: sum|q| often > 32767
        packed = orc.pack(type_, w)
        ref = orc.mul_mat(type_, packed, x, N)
        for rep in range(2):
            out = run(ctx, type_, packed, x, N, a.ALGO_TCS)
            bad = bits(out) != bits(ref)
            assert not bad.any(), f"type={type_} K={K} rep={rep}: {bad.sum()} of {bad.size} differ, first at {np.argwhere(bad)[0]}"


@pytest.mark.parametrize("T", [5, 32, 48])
def test_gemm_on_a_workspace_quantised_for_another_kernel(ctx, orc, T):
    """vlut_quantize_act lays the int8 operand tiles out for the kernel AUTO would run (tile height, and for the small-T
    kernel the regrouped K order); a vlut_gemm that names another kernel must re-tile the workspace first, both ways."""
    a = api()
    type_, K, N = O.I1_V_2, 3200, 264
    w = O.make_ternary(N, K, seed=21)
    x = O.make_acts(T, K, seed=T)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    W = ctx.upload(type_, K, N, packed)
    xd = torch.from_numpy(x).cuda()
    other = a.ALGO_TCP if T > 32 else a.ALGO_SIMPLE
    for algo in [other, a.ALGO_TCS, a.ALGO_SIMPLE, a.ALGO_TCS, a.ALGO_AUTO]:
        ctx.quantize_act(type_, xd)
        out = ctx.gemm(W, T, algo=algo)
        torch.cuda.synchronize()
        assert np.array_equal(bits(out.cpu().numpy()), bits(ref)), f"algo={algo}"
        out2 = ctx.gemm(W, T, algo=other)   # and back, on the same quantisation
        torch.cuda.synchronize()
        assert np.array_equal(bits(out2.cpu().numpy()), bits(ref)), f"algo={algo} then {other}"


def test_fused_scatter_epilogue_single_gpu_emulation(ctx, orc):
    """The fused all-gather epilogue (vlut_mul_mat_scatter, peer-pointer flavour), with the 'ranks' emulated as two
    shards and two result buffers on ONE GPU: each shard's GEMM writes its column slice into both [T][N] buffers."""
    a = api()
    type_, K, N, T = O.I2_V_4, 2048, 640, 40
    w = O.make_ternary(N, K, seed=3)
    x = O.make_acts(T, K, seed=4)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    xd = torch.from_numpy(x).cuda()
    bufs = [torch.zeros((T, N), device="cuda") for _ in range(2)]
    bounds = [0, 384, 640]
    for g in range(2):
        r0, rows = bounds[g], bounds[g + 1] - bounds[g]
        Wg = ctx.upload(type_, K, rows, a.shard_rows(type_, K, N, r0, rows, packed))
        ctx.mul_mat_scatter(Wg, xd, peers=[b.data_ptr() for b in bufs], multicast_ptr=0, ldo=N, col0=r0)
    torch.cuda.synchronize()
    for b in bufs:
        assert np.array_equal(bits(b.cpu().numpy()), bits(ref))


@pytest.mark.parametrize("T", [4, 16, 64])
def test_inputs_visible_with_an_unrelated_matmul_in_between(orc, T):
    """Programmatic-launch chains nest: in G1(->y1), Q2, G2(->y3), Q3(x = y1) the early-reading Q3 only orders itself after
    G2's CTAs have STARTED, so it must not read y1 ahead of its dependency wait while G1 may still be storing it.  The
    library remembers the output ranges of every GEMM in flight (not just the last one).  Small shape: G1 and G2 fit on
    the GPU together."""
    a = api()
    c = a.Context(0)
    c.set_option(a.Context.OPT_INPUTS_VISIBLE, 1)
    type_, K = O.I2_V_4, 1024
    packed = [orc.pack(type_, O.make_ternary(K, K, seed=s)) for s in (5, 6, 7)]
    W = [c.upload(type_, K, K, p) for p in packed]
    x0, x1 = O.make_acts(T, K, seed=8), O.make_acts(T, K, seed=9)
    xd0, xd1 = torch.from_numpy(x0).cuda(), torch.from_numpy(x1).cuda()
    y1, y2, y3 = (torch.empty((T, K), device="cuda") for _ in range(3))
    r1 = orc.mul_mat(type_, packed[0], x0, K)
    r3 = orc.mul_mat(type_, packed[1], x1, K)
    r2 = orc.mul_mat(type_, packed[2], r1, K)
    for rep in range(20):
        y1.zero_()
        torch.cuda.synchronize()
        c.mul_mat(W[0], xd0, out=y1)      # G1 -> y1
        c.mul_mat(W[1], xd1, out=y3)      # unrelated: Q2 (early), G2 -> y3
        c.mul_mat(W[2], y1, out=y2)       # Q3 reads y1: early mode must be refused
        torch.cuda.synchronize()
        assert np.array_equal(bits(y2.cpu().numpy()), bits(r2)), f"rep {rep}"
        assert np.array_equal(bits(y3.cpu().numpy()), bits(r3)), f"rep {rep}"
    c.close()


@pytest.mark.parametrize("T", [1, 3, 16, 40, 64, 200])
@pytest.mark.parametrize("type_,K,Ns", [(O.I2_V_4, 4096, [256, 128, 128]), (O.I1_V_2, 3200, [520, 520]),
                                        (O.I2_V, 1040, [129, 77, 300, 64, 130])])
def test_mul_mat_batch_equals_separate_ops(ctx, orc, type_, K, Ns, T):
    """vlut_mul_mat_batch: several weights on ONE activation tensor (Q/K/V, up/gate) -- one quantisation and, at T <= 64,
    one launch of the streaming kernel over all strips -- is bit-identical to separate MUL_MATs (and to the oracle)."""
    x = O.make_acts(T, K, seed=T, kind="uniform" if T % 2 else "normal")
    xd = torch.from_numpy(x).cuda()
    packed = [orc.pack(type_, O.make_ternary(n, K, seed=n)) for n in Ns]
    W = [ctx.upload(type_, K, n, p) for n, p in zip(Ns, packed)]
    scales = [1.0, 0.5, 1.25, 1.0, 3.0][: len(Ns)]
    l0 = ctx.launch_count
    outs = ctx.mul_mat_batch(W, xd, w_scales=scales)
    n_launch = ctx.launch_count - l0
    torch.cuda.synchronize()
    for n, p, o, sc in zip(Ns, packed, outs, scales):
        ref = orc.mul_mat(type_, p, x, n)
        if sc != 1.0:
            ref = ref * np.float32(sc)
        assert np.array_equal(bits(o.cpu().numpy()), bits(ref)), f"N={n}"
    if T <= 64 and len(Ns) <= 4:
        assert n_launch <= 2, f"{n_launch} launches for one batch"   # quantise + one GEMM (or one fused kernel)


@pytest.mark.parametrize("split", [0, 1, 2, 5])
@pytest.mark.parametrize("T", [1, 2, 3, 4, 7, 8])
def test_gemv_fused_quantise_kernel(ctx, orc, knobs, split, T):
    """VLUT_ALGO_GEMV: the whole op in one launch on the CUDA cores (quantisation fused, K split over a cluster).  Forced
    cluster sizes, ragged N, I1 tails (K % 20 != 0, K % 640 != 0), strided x, batch of weights."""
    a = api()
    knobs(VLUT_GEMV_SPLIT=split)
    for type_, K, N in [(O.I2_V_4, 4096, 300), (O.I1_V_2, 3200, 130), (O.I2_V, 1040, 129), (O.I1_V, 3072, 72),
                        (O.I2_V_4, 14336, 128), (O.I1_V_2, 8640, 64)]:
        w = O.make_ternary(N, K, seed=K + split)
        x = O.make_acts(T, K, seed=T, kind="uniform" if T % 2 else "outlier")
        packed = orc.pack(type_, w)
        ref = orc.mul_mat(type_, packed, x, N)
        W = ctx.upload(type_, K, N, packed)
        buf = torch.zeros((T, K + 64), device="cuda")
        buf[:, :K] = torch.from_numpy(x).cuda()
        out = ctx.mul_mat(W, buf[:, :K], algo=a.ALGO_GEMV)
        torch.cuda.synchronize()
        assert a.ALGO_NAMES[ctx.last_algo] == "gemv"
        bad = bits(out.cpu().numpy()) != bits(ref)
        assert not bad.any(), f"type={type_} K={K}: {bad.sum()} of {bad.size} differ, first at {np.argwhere(bad)[0]}"
    # int16 wrap inside every K-group
    K, N = 2048, 144
    wv = np.ones((N, K), dtype=np.float32)
    wv[::3] = -1.0
    xv = O.make_acts(T, K, kind="ones")
    xv[1::2] *= -1.0
    pk = orc.pack(O.I2_V_4, wv)
    out = ctx.mul_mat(ctx.upload(O.I2_V_4, K, N, pk), torch.from_numpy(xv).cuda(), algo=a.ALGO_GEMV)
    torch.cuda.synchronize()
    assert np.array_equal(bits(out.cpu().numpy()), bits(orc.mul_mat(O.I2_V_4, pk, xv, N)))


def test_op_supported_mirrors_the_launch_path(ctx):
    a = api()
    assert a.op_supported(O.I2_V_4, 4096, 14336, 512)
    assert a.op_supported(O.I1_V_2, 3200, 8640, 1)
    assert not a.op_supported(O.I2_V_4, 4104, 16, 1)          # K % 16 != 0 (the reference packer would overrun)
    assert not a.op_supported(O.I2_V_4, 65536, 16, 1)         # more than 64 K-groups
    assert not a.op_supported(7, 4096, 16, 1)                 # not a vlut type
    assert a.op_supported(O.I2_V_4, 32768, 16, 1)


def test_barrier_kernel_two_ranks_emulated_on_one_gpu():
    """vlut_barrier with the two 'ranks' as two contexts (two streams) on ONE GPU: each waits for the other's flag; the
    epoch lives on the device, so repeated calls (and graph replays) stay in step."""
    a = api()
    s0, s1 = torch.cuda.Stream(), torch.cuda.Stream()
    c0, c1 = a.Context(0, stream=s0), a.Context(0, stream=s1)
    flags = [torch.zeros(16, dtype=torch.int32, device="cuda") for _ in range(2)]
    ptrs = [f.data_ptr() for f in flags]
    torch.cuda.synchronize()
    for rep in range(5):
        c0.barrier(ptrs, 0)
        c1.barrier(ptrs, 1)
    torch.cuda.synchronize()
    for r, f in enumerate(flags):
        h = f.cpu().numpy()
        assert h[0] == 5 and h[1] == 5 and h[8] == 5, (r, h)
    c0.close()
    c1.close()


@pytest.mark.parametrize("vec", [1, 0])
@pytest.mark.parametrize("T,N,bounds", [(40, 640, [0, 384, 640]), (130, 1000, [0, 512, 1000]), (16, 262, [0, 128, 262]),
                                        (200, 390, [0, 130, 390])])
def test_fused_scatter_vector_stores(ctx, orc, knobs, vec, T, N, bounds):
    """The scatter epilogue with 16-byte stores (four rows per lane after a 4 x 4 lane transpose): aligned shards, ragged
    row tails (N % 4 != 0), a shard whose first column is not 16-byte aligned (scalar fallback), and VLUT_SCATTER_VEC=0."""
    a = api()
    knobs(VLUT_SCATTER_VEC=vec)
    type_, K = O.I2_V_4, 2048
    w = O.make_ternary(N, K, seed=N)
    x = O.make_acts(T, K, seed=T)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    xd = torch.from_numpy(x).cuda()
    ld = (N + 3) // 4 * 4
    bufs = [torch.zeros((T, ld), device="cuda") for _ in range(2)]
    for g in range(2):
        r0, rows = bounds[g], bounds[g + 1] - bounds[g]
        Wg = ctx.upload(type_, K, rows, a.shard_rows(type_, K, N, r0, rows, packed))
        ctx.mul_mat_scatter(Wg, xd, peers=[b.data_ptr() for b in bufs], multicast_ptr=0, ldo=ld, col0=r0)
    torch.cuda.synchronize()
    for b in bufs:
        assert np.array_equal(bits(b[:, :N].cpu().numpy()), bits(ref))
        assert (b[:, N:] == 0).all()


@pytest.mark.parametrize("bulk", [1, 0], ids=["bulk-copies", "direct-stores"])
@pytest.mark.parametrize("T,N,bounds,K,kind,extra", [
    (300, 1536, [0, 768, 1536], 2048, "normal", {}),
    (520, 1000, [0, 512, 1000], 4096, "normal", {}),                   # ragged last strip (488 rows: 416-byte copies), ragged T
    (300, 1536, [0, 768, 1536], 2048, "ones", {}),                     # int16 wrap: exact path (direct stores)
    (300, 1536, [0, 640, 1536], 4096, "normal", {"VLUT_TC_GRID": 6}),  # split tiles: the last arriver's fix-up scatters
    (520, 1536, [0, 512, 1024, 1536], 2048, "outlier", {"VLUT_TC_NT": 256}),   # 256-token tile, three destinations
], ids=["plain", "ragged", "exact-path", "split-tiles", "nt256-3way"])
def test_fused_scatter_bulk_copies(ctx, orc, knobs, bulk, T, N, bounds, K, kind, extra):
    """CTA-pair kernel, peer-pointer scatter through shared-memory staging + cp.async.bulk (epilogue.cuh, Scatter) against the
    direct-store scatter and the oracle: every 'rank' buffer ends up with the whole [T][N] result, bit for bit."""
    a = api()
    knobs(VLUT_SCATTER_BULK=bulk, **extra)
    type_ = O.I2_V_4
    w = O.make_ternary(N, K, seed=N + 1) if kind != "ones" else np.ones((N, K), dtype=np.float32)
    x = O.make_acts(T, K, seed=T, kind=kind)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    xd = torch.from_numpy(x).cuda()
    G = len(bounds) - 1
    bufs = [torch.full((T + 1, N + 4), -7.0, device="cuda") for _ in range(G)]   # guard row and guard columns
    for g in range(G):
        r0, rows = bounds[g], bounds[g + 1] - bounds[g]
        Wg = ctx.upload(type_, K, rows, a.shard_rows(type_, K, N, r0, rows, packed))
        for rep in range(2):   # twice: the staging buffers and the copy groups are reused across launches
            ctx.mul_mat_scatter(Wg, xd, peers=[b.data_ptr() for b in bufs], multicast_ptr=0, ldo=N + 4, col0=r0)
    torch.cuda.synchronize()
    for b in bufs:
        assert np.array_equal(bits(b[:T, :N].cpu().numpy()), bits(ref))
        assert (b[T:] == -7.0).all() and (b[:, N:] == -7.0).all()


@pytest.mark.parametrize("fast", [1, 0], ids=["fastpath-allowed", "per-group-drain"])
@pytest.mark.parametrize("flat", [1, 0], ids=["flat-walk-allowed", "k-group-units"])
@pytest.mark.parametrize("lean,astages,bulk_out,nt", [(1, 2, 0, 128), (0, 2, 0, 128), (1, 3, 0, 128), (1, 4, 0, 128),
                                                     (1, 2, 1, 128), (0, 4, 1, 128), (1, 4, 0, 256), (0, 2, 1, 256)])
@pytest.mark.parametrize("T", [65, 130, 300, 700])
def test_tcp_role_loop_and_drain_variants(ctx, orc, knobs, lean, astages, bulk_out, nt, T, fast, flat):
    """Every shipped variant of the CTA-pair kernel's role loops and drain: lean loops (uniform shapes) vs the generic
    unit-walking loops (VLUT_TC_LEAN=0, and always for the I1 types' 640-k groups and K not a multiple of the group size),
    two / three (single accumulator) / four A stages, both token tiles, direct stores vs the staged drain through one 2-D TMA
    tensor store per chunk (VLUT_TC_BULK_OUT=1; needs 16-byte aligned rows: N % 4 == 0, the ragged N = 900 shape clips at
    the tensor bound), fast and exact (per-K-group) drains, and the flat walk (K in uniform A blocks across the K-group
    boundaries when the quantise kernel reports that no token tile can wrap: normal activations take it, uniform ones
    do not) for the shapes whose groups do not fit the lean loops -- 640-k groups with a short last block (K = 3200), with
    whole blocks (K = 8640 -> 8704), K not a multiple of the group size (K = 1040)."""
    a = api()
    knobs(VLUT_TC_LEAN=lean, VLUT_TC_ASTAGES=astages, VLUT_TC_BULK_OUT=bulk_out, VLUT_TC_NT=nt, VLUT_TC_FAST=fast,
          VLUT_TC_FLAT=flat)
    for type_, K, N in [(O.I2_V_4, 2048, 1024), (O.I2_V_4, 4096, 900), (O.I1_V_2, 3200, 520), (O.I2_V, 1040, 388),
                        (O.I1_V_2, 8640, 260)]:
        w = O.make_ternary(N, K, seed=T + astages)
        x = O.make_acts(T, K, seed=T, kind="uniform" if (T + astages) % 2 else "normal")
        packed = orc.pack(type_, w)
        ref = orc.mul_mat(type_, packed, x, N)
        for rep in range(2):
            out = run(ctx, type_, packed, x, N, a.ALGO_TCP)
            bad = bits(out) != bits(ref)
            assert not bad.any(), f"{O.TYPE_NAMES[type_]} K={K} N={N} rep={rep}: {bad.sum()} of {bad.size} differ"


@pytest.mark.parametrize("rows", [128, 112, 100, 64])
@pytest.mark.parametrize("split", [1, 2])
@pytest.mark.parametrize("T", [1, 16, 48])
def test_tcs_virtual_strips(ctx, orc, knobs, rows, split, T):
    """VLUT_TCS_ROWS: the small-T kernel with fewer than 128 weight rows per CTA (a CTA's rows then straddle two T2 strips,
    lanes beyond its rows fetch and store nothing), alone and combined with the cluster split; single tensors with ragged N
    and the batched launch (several tensors in one strip index space)."""
    a = api()
    knobs(VLUT_TCS_ROWS=rows, VLUT_TCS_SPLIT=split)
    for type_, K, N in [(O.I2_V_4, 2048, 1000), (O.I1_V_2, 3200, 520), (O.I2_V, 1040, 129)]:
        w = O.make_ternary(N, K, seed=rows)
        x = O.make_acts(T, K, seed=T)
        packed = orc.pack(type_, w)
        ref = orc.mul_mat(type_, packed, x, N)
        out = run(ctx, type_, packed, x, N, a.ALGO_TCS)
        assert np.array_equal(bits(out), bits(ref)), f"{O.TYPE_NAMES[type_]} K={K} N={N}"
    # batched: three tensors of different heights share the activation
    K, Ns = 2048, [520, 128, 260]
    x = O.make_acts(T, K, seed=3)
    packs = [orc.pack(O.I2_V_4, O.make_ternary(n, K, seed=n)) for n in Ns]
    ws = [ctx.upload(O.I2_V_4, K, n, p) for n, p in zip(Ns, packs)]
    outs = ctx.mul_mat_batch(ws, torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    for n, p, o in zip(Ns, packs, outs):
        assert np.array_equal(bits(o.cpu().numpy()), bits(orc.mul_mat(O.I2_V_4, p, x, n))), f"batched N={n}"


@pytest.mark.parametrize("T,kind", [(40, "normal"), (130, "normal"), (300, "uniform"), (520, "normal")])
@pytest.mark.parametrize("type_,K,N,bounds", [(O.I2_V_4, 2048, 640, [0, 384, 640]), (O.I1_V_2, 3200, 900, [0, 512, 900]),
                                              (O.I2_V_4, 4096, 1800, [0, 640, 1280, 1800])])
def test_scatter_push_single_gpu_emulation(ctx, orc, type_, K, N, bounds, T, kind):
    """vlut_mul_mat_scatter_push (local GEMM that stamps its tiles + the copy kernel on a second stream that forwards them),
    the 'ranks' emulated as shards and result buffers on ONE GPU: every shard's call writes its column slice into its own
    buffer and the copy kernel forwards it to the others.  T <= 64 takes the documented fall-back (fused scatter); uniform
    activations the exact path; split tiles (hybrid) stamp from the last arriver; twice, so that the epoch-stamped tile
    flags of the first call are stale for the second."""
    a = api()
    w = O.make_ternary(N, K, seed=5)
    x = O.make_acts(T, K, seed=6, kind=kind)
    packed = orc.pack(type_, w)
    ref = orc.mul_mat(type_, packed, x, N)
    xd = torch.from_numpy(x).cuda()
    G = len(bounds) - 1
    for rep in range(2):
        bufs = [torch.zeros((T, N), device="cuda") for _ in range(G)]
        for g in range(G):
            r0, rows = bounds[g], bounds[g + 1] - bounds[g]
            Wg = ctx.upload(type_, K, rows, a.shard_rows(type_, K, N, r0, rows, packed))
            ctx.mul_mat_scatter_push(Wg, xd, peers=[b.data_ptr() for b in bufs], self_index=g, ldo=N, col0=r0)
        torch.cuda.synchronize()
        for i, b in enumerate(bufs):
            assert np.array_equal(bits(b.cpu().numpy()), bits(ref)), f"rep {rep} buffer {i}"
