#!/usr/bin/env python
"""Write a random-init Llama-architecture GGUF with TERNARY projection weights, for driving the reference's own tools
(llama-quantize, llama-bench, llama-batched-bench) on the B200 plugin.  No checkpoint, no network: synthetic weights of
the named model's shapes, tokenizer metadata copied from the reference's vocab-only file
(baseline/_ref/models/ggml-vocab-llama-spm.gguf -- data, not source).

Two modes:
  --ftype f32      every tensor f32 (ternary values {-1,0,+1} for the seven projection matrices of a layer, plus their {1}
                   `.scale` tensors): the input of the STOCK `llama-quantize ... I2_V_4` (kept small: --layers 2)
  --ftype i2_v_4 / i2_v / i2_v_8 / i1_v / i1_v_2
                   the projection matrices are written already packed (uniform bytes in [0,81) resp. [0,243) ARE iid
                   uniform ternary digits, i.e. a valid packing -- the same trick bench.py uses), token embedding and output
                   in Q8_0: the full 32-layer model without a 30 GB f32 detour.  One layer's bytes are reused for all
                   layers (synthetic timing does not care).

A minimal GGUF v3 writer (header, KV section, tensor infos, 32-byte aligned data) -- the format is documented in
ggml/include/gguf.h of the reference; nothing here is copied from gguf-py.
"""
from __future__ import annotations

import argparse
import struct
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ALIGN = 32
T_U8, T_I8, T_U16, T_I16, T_U32, T_I32, T_F32, T_BOOL, T_STR, T_ARR, T_U64, T_I64, T_F64 = range(13)
SCALAR_SIZE = {T_U8: 1, T_I8: 1, T_U16: 2, T_I16: 2, T_U32: 4, T_I32: 4, T_F32: 4, T_BOOL: 1, T_U64: 8, T_I64: 8, T_F64: 8}
GGML_F32, GGML_Q8_0 = 0, 8
VLUT_TYPES = {"i2_v": 40, "i2_v_4": 42, "i2_v_8": 43, "i1_v": 45, "i1_v_2": 46}
FTYPE_ID = {"f32": 0, "i2_v": 38, "i2_v_4": 39, "i2_v_8": 40, "i1_v": 41, "i1_v_2": 42}

SHAPES = {  # name: (n_embd, n_ff, n_head, n_head_kv)
    "llama3-8b": (4096, 14336, 32, 8),
    "bitnet-3b": (3200, 8640, 32, 32),
    "falcon3-1b": (2048, 8192, 8, 4),
    "trilm-3.9b": (3072, 9216, 24, 24),
    "tiny": (512, 1536, 8, 4),   # test-sized
}


def gstr(s: str | bytes) -> bytes:
    b = s if isinstance(s, bytes) else s.encode("utf-8")
    return struct.pack("<Q", len(b)) + b


def read_tokenizer_kvs(path: Path) -> list[bytes]:
    """The raw bytes of every `tokenizer.*` key/value pair of a GGUF file (re-emitted verbatim)."""
    data = path.read_bytes()
    magic, ver, nt, nkv = struct.unpack_from("<IIQQ", data, 0)
    assert magic == 0x46554747 and ver == 3, "not a GGUF v3 file"
    off = 24
    out = []

    def skip_str(o):
        n, = struct.unpack_from("<Q", data, o)
        return o + 8 + n

    def skip_val(t, o):
        if t == T_STR:
            return skip_str(o)
        if t == T_ARR:
            et, cnt = struct.unpack_from("<IQ", data, o)
            o += 12
            if et == T_STR:
                for _ in range(cnt):
                    o = skip_str(o)
                return o
            return o + SCALAR_SIZE[et] * cnt
        return o + SCALAR_SIZE[t]

    for _ in range(nkv):
        start = off
        n, = struct.unpack_from("<Q", data, off)
        key = data[off + 8: off + 8 + n].decode()
        off += 8 + n
        t, = struct.unpack_from("<I", data, off)
        off = skip_val(t, off + 4)
        if key.startswith("tokenizer."):
            out.append(data[start:off])
    return out


def kv(key: str, typ: int, val) -> bytes:
    b = gstr(key) + struct.pack("<I", typ)
    if typ == T_STR:
        return b + gstr(val)
    fmt = {T_U32: "<I", T_F32: "<f", T_BOOL: "<B", T_U64: "<Q", T_I32: "<i"}[typ]
    return b + struct.pack(fmt, val)


def q8_0(x: np.ndarray) -> np.ndarray:
    """f32 [rows][cols] -> Q8_0 bytes (blocks of 32: f16 d, 32 x int8), cols % 32 == 0."""
    r, c = x.shape
    blk = x.reshape(r, c // 32, 32)
    d = np.abs(blk).max(axis=2) / 127.0
    inv = np.where(d > 0, 1.0 / np.where(d > 0, d, 1), 0).astype(np.float32)
    q = np.rint(blk * inv[:, :, None]).astype(np.int8)
    out = np.empty((r, c // 32, 34), dtype=np.uint8)
    out[:, :, :2] = d.astype(np.float16).view(np.uint8).reshape(r, c // 32, 2)
    out[:, :, 2:] = q.view(np.uint8)
    return out.reshape(-1)


def vlut_row_bytes(ftype: str, K: int) -> int:
    return (K // 20 * 4 + K % 20 // 4) if ftype.startswith("i1") else K // 4


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", required=True)
    ap.add_argument("--model", default="llama3-8b", choices=sorted(SHAPES))
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--ftype", default="f32", choices=sorted(FTYPE_ID))
    ap.add_argument("--vocab", default=str(ROOT / "baseline" / "_ref" / "models" / "ggml-vocab-llama-spm.gguf"))
    ap.add_argument("--ctx", type=int, default=8192)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--no-scale", action="store_true", help="omit the {1} `.scale` tensors")
    args = ap.parse_args()

    n_embd, n_ff, n_head, n_head_kv = SHAPES[args.model]
    head_dim = n_embd // n_head
    n_kv = n_head_kv * head_dim
    n_vocab = 32000
    rng = np.random.Generator(np.random.PCG64(args.seed))
    packed = args.ftype != "f32"

    kvs = [
        kv("general.architecture", T_STR, "llama"),
        kv("general.name", T_STR, f"synthetic-ternary-{args.model}-{args.layers}L"),
        kv("general.file_type", T_U32, FTYPE_ID[args.ftype]),
        kv("llama.block_count", T_U32, args.layers),
        kv("llama.context_length", T_U32, args.ctx),
        kv("llama.embedding_length", T_U32, n_embd),
        kv("llama.feed_forward_length", T_U32, n_ff),
        kv("llama.attention.head_count", T_U32, n_head),
        kv("llama.attention.head_count_kv", T_U32, n_head_kv),
        kv("llama.attention.layer_norm_rms_epsilon", T_F32, 1e-5),
        kv("llama.rope.dimension_count", T_U32, head_dim),
        kv("llama.rope.freq_base", T_F32, 500000.0),
        kv("llama.vocab_size", T_U32, n_vocab),
    ] + read_tokenizer_kvs(Path(args.vocab))

    # tensors: (name, ggml dims (ne0 fastest), ggml type, producer of the bytes)
    tensors = []

    def add(name, dims, typ, make):
        tensors.append((name, dims, typ, make))

    def dense_f32(rows, cols, std):
        return lambda: (rng.standard_normal((rows, cols), dtype=np.float32) * np.float32(std)).view(np.uint8).reshape(-1)

    def emb(rows, cols):
        if packed:
            return GGML_Q8_0, (lambda: q8_0(rng.standard_normal((rows, cols), dtype=np.float32) * np.float32(0.05)))
        return GGML_F32, dense_f32(rows, cols, 0.05)

    cache: dict = {}

    def ternary(tag, K, N):
        """[N][K] ternary matrix: packed vlut bytes (reused across layers) or f32 {-1,0,+1}."""
        if packed:
            def make():
                if (tag, K, N) not in cache:
                    hi = 243 if args.ftype.startswith("i1") else 81
                    nb = vlut_row_bytes(args.ftype, K) * N
                    b = rng.integers(0, hi, size=nb, dtype=np.uint8)
                    if args.ftype.startswith("i1") and K % 20:   # the 4-trit tail bytes hold values < 81
                        nb5 = K // 20 * 4 * N
                        b[nb5:] = rng.integers(0, 81, size=nb - nb5, dtype=np.uint8)
                    cache[(tag, K, N)] = b
                return cache[(tag, K, N)]
            return VLUT_TYPES[args.ftype], make
        return GGML_F32, (lambda: (rng.integers(0, 3, size=(N, K), dtype=np.int8) - 1).astype(np.float32).view(np.uint8).reshape(-1))

    t, m = emb(n_vocab, n_embd)
    add("token_embd.weight", [n_embd, n_vocab], t, m)
    ones = lambda n: (lambda: np.ones(n, dtype=np.float32).view(np.uint8))  # noqa: E731
    for il in range(args.layers):
        p = f"blk.{il}."
        add(p + "attn_norm.weight", [n_embd], GGML_F32, ones(n_embd))
        for nm, K, N in [("attn_q", n_embd, n_embd), ("attn_k", n_embd, n_kv), ("attn_v", n_embd, n_kv),
                         ("attn_output", n_embd, n_embd)]:
            t, m = ternary(nm, K, N)
            add(p + nm + ".weight", [K, N], t, m)
            if not args.no_scale:
                # keeps activations O(1): a ternary row of K entries has norm ~ sqrt(2K/3)
                sc = float(1.0 / np.sqrt(2.0 * K / 3.0))
                add(p + nm + ".scale", [1], GGML_F32, (lambda s=sc: np.array([s], dtype=np.float32).view(np.uint8)))
        add(p + "ffn_norm.weight", [n_embd], GGML_F32, ones(n_embd))
        for nm, K, N in [("ffn_gate", n_embd, n_ff), ("ffn_up", n_embd, n_ff), ("ffn_down", n_ff, n_embd)]:
            # (the reference's Llama loader creates `.scale` tensors for the attention projections only,
            # src/llama-model.cpp:2721-2728; an FFN scale would be an unconsumed tensor and fail the load)
            t, m = ternary(nm, K, N)
            add(p + nm + ".weight", [K, N], t, m)
    add("output_norm.weight", [n_embd], GGML_F32, ones(n_embd))
    t, m = emb(n_vocab, n_embd)
    add("output.weight", [n_embd, n_vocab], t, m)

    def nbytes(dims, typ):
        n = int(np.prod(dims))
        if typ == GGML_F32:
            return 4 * n
        if typ == GGML_Q8_0:
            return n // 32 * 34
        inv = {v: k for k, v in VLUT_TYPES.items()}
        return vlut_row_bytes(inv[typ], dims[0]) * dims[1]

    infos, off = b"", 0
    offsets = []
    for name, dims, typ, _ in tensors:
        infos += gstr(name) + struct.pack("<I", len(dims)) + b"".join(struct.pack("<Q", d) for d in dims)
        infos += struct.pack("<IQ", typ, off)
        offsets.append(off)
        off += -(-nbytes(dims, typ) // ALIGN) * ALIGN

    out = Path(args.out)
    out.parent.mkdir(parents=True, exist_ok=True)
    with out.open("wb") as f:
        f.write(struct.pack("<IIQQ", 0x46554747, 3, len(tensors), len(kvs)))
        for b in kvs:
            f.write(b)
        f.write(infos)
        pad = -f.tell() % ALIGN
        f.write(b"\0" * pad)
        base = f.tell()
        for (name, dims, typ, make), o in zip(tensors, offsets):
            assert f.tell() == base + o, name
            b = make()
            assert b.size == nbytes(dims, typ), (name, b.size, nbytes(dims, typ))
            f.write(b.tobytes())
            f.write(b"\0" * (-b.size % ALIGN))
    print(f"wrote {out} ({out.stat().st_size / 1e6:.1f} MB): {args.model}, {args.layers} layers, {args.ftype}, "
          f"{len(tensors)} tensors")


if __name__ == "__main__":
    main()
