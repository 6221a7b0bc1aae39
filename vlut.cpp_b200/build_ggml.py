"""Build the ggml backend plugin libggml-vlut-b200.so against the reference's own (unmodified) ggml headers.

The headers are read where they lie (never copied): $VLUT_GGML_ROOT or /root/reference/ggml.  Where they are not
reachable (the GPU box) the prebuilt plugin that travelled with the snapshot is kept.
"""
from __future__ import annotations

import os
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = HERE / "csrc" / "ggml_backend" / "ggml-vlut-b200.cpp"
OUT = HERE / "csrc" / "build"
PLUGIN = OUT / "libggml-vlut-b200.so"


def ggml_root() -> Path | None:
    for cand in (os.environ.get("VLUT_GGML_ROOT"), "/root/reference/ggml"):
        if cand and (Path(cand) / "src" / "ggml-backend-impl.h").exists():
            return Path(cand)
    return None


def build(force: bool = False) -> dict:
    root = ggml_root()
    if root is None:
        return {"ggml_plugin": str(PLUGIN) if PLUGIN.exists() else "skipped (ggml headers not reachable)"}
    deps = [SRC, HERE.parent / "include" / "vlut_b200.h", OUT / "libvlut_b200.so"]
    if not force and PLUGIN.exists() and all(PLUGIN.stat().st_mtime >= d.stat().st_mtime for d in deps if d.exists()):
        return {"ggml_plugin": str(PLUGIN)}
    gxx = "/usr/bin/g++" if Path("/usr/bin/g++").exists() else "g++"
    cmd = [gxx, "-O2", "-fPIC", "-shared", "-std=c++17", "-DGGML_BACKEND_DL", "-DGGML_BACKEND_SHARED", "-DGGML_BACKEND_BUILD",
           "-DGGML_SHARED", f"-I{root}/include", f"-I{root}/src", str(SRC), "-o", str(PLUGIN), f"-L{OUT}", "-lvlut_b200",
           "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"building the ggml plugin failed:\n{r.stdout}")
    return {"ggml_plugin": str(PLUGIN)}
