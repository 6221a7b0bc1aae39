"""Build the sm_100a CUDA library (and, when the ggml headers are reachable, the ggml backend plugin) in-tree.

Everything is compiled with explicit `nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`; the resulting
`.so` files live under vlut.cpp_b200/csrc/build/ (git-ignored, shipped to the GPU box by gpurun).
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
OUT = CSRC / "build"
LIB = OUT / "libvlut_b200.so"
GGML_PLUGIN = OUT / "libggml-vlut-b200.so"

CU_SOURCES = ["c_api.cu", "repack.cu", "quantize.cu", "gemm_simple.cu", "gemm_tc5.cu", "gemm_tc5_nt64.cu", "gemm_tc5_nt128.cu", "gemm_tc5_nt128b.cu", "gemm_tc5_nt128c.cu", "gemm_tc5_nt256.cu", "gemm_tc5_nt256b.cu", "gemm_tcs.cu", "gemm_gemv.cu", "gemm_lut.cu", "peak.cu", "barrier.cu", "push.cu", "group.cu", "no_alloc.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC"]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found: the CUDA library cannot be built")


def _stale(target: Path, deps: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps if d.exists())


def build_lib(force: bool = False, verbose: bool = False) -> Path:
    OUT.mkdir(parents=True, exist_ok=True)
    srcs = [CSRC / s for s in CU_SOURCES]
    deps = srcs + list(CSRC.glob("*.cuh")) + [HERE.parent / "include" / "vlut_b200.h"]
    if not force and not _stale(LIB, deps):
        return LIB
    objs = []
    procs = []
    for s in srcs:
        o = OUT / (s.stem + ".o")
        objs.append(o)
        if force or _stale(o, [s] + list(CSRC.glob("*.cuh")) + [HERE.parent / "include" / "vlut_b200.h"]):
            cmd = [nvcc(), *NVCC_FLAGS, "-c", str(s), "-o", str(o)]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {s.name}:\n{out}")
        if verbose and out.strip():
            print(out)
    cmd = [nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", str(LIB), *map(str, objs)]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc link failed:\n{r.stdout}")
    return LIB


def build_all(force: bool = False, verbose: bool = False) -> dict:
    built = {"lib": str(build_lib(force=force, verbose=verbose))}
    try:
        from . import build_ggml  # package import (vlut_b200.build)
    except ImportError:  # run as a script
        import importlib.util
        spec = importlib.util.spec_from_file_location("build_ggml", HERE / "build_ggml.py")
        build_ggml = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(build_ggml)
    built.update(build_ggml.build(force=force))
    return built


if __name__ == "__main__":
    import sys
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
