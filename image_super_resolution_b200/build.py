"""Build libsrb200.so in-tree with nvcc for sm_100a (no torch extension machinery, plain C ABI)."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "srb200.cu")
OUT = os.path.join(HERE, "libsrb200.so")
OUT_DEBUG = os.path.join(HERE, "libsrb200_dbg.so")     # -DSRB_DEBUG: probe hooks for tools/ (SRB200_LIB=...); never loaded by the product
DEPS = sorted(os.path.join(HERE, "csrc", f) for f in os.listdir(os.path.join(HERE, "csrc"))) + [
    os.path.join(HERE, "..", "include", "srb200.h"), os.path.join(HERE, "..", "include", "srb200_debug.h")]


def nvcc_path() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def up_to_date(out: str = OUT) -> bool:
    if not os.path.exists(out):
        return False
    t = os.path.getmtime(out)
    return all(os.path.getmtime(d) <= t for d in DEPS if os.path.exists(d))


def build(force: bool = False, verbose: bool = False, debug: bool = False) -> str:
    out = OUT_DEBUG if debug else OUT
    if up_to_date(out) and not force:
        return out
    cmd = [nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "-Xcompiler", "-fPIC", "-shared", "-o", out, SRC]
    if debug:
        cmd[1:1] = ["-DSRB_DEBUG"]
    if verbose:
        cmd[1:1] = ["-Xptxas", "-v"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return out


if __name__ == "__main__":
    import sys
    print(build(force=True, verbose=True, debug="--debug" in sys.argv))
