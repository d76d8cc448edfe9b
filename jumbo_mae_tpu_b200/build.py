"""Builds libjumbo_mae_b200.so (sm_100a only) in-tree with nvcc.

Usage: python -m jumbo_mae_tpu_b200.build [--force]
The library is written next to this file so that it travels with the repo snapshot
to the GPU box; it is git-ignored.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libjumbo_mae_b200.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr",
] + (["-DJM_DEBUG_SMID"] if os.environ.get("JM_DEBUG_SMID") else []) + \
    (["-DJM_EXP_NO_STORE"] if os.environ.get("JM_EXP_NO_STORE") else []) + \
    os.environ.get("JM_NVCC_DEFS", "").split()          # experiments: extra -D flags


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or "/usr/local/cuda/bin/nvcc"
    return cand if os.path.exists(cand) else "nvcc"


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _flag_stamp() -> str:
    """The compile flags an object was built with are written next to it (``<obj>.flags``): an object built with
    other -D switches (JM_NVCC_DEFS, JM_DEBUG_SMID, ...) is stale even when its sources did not change."""
    return " ".join([_nvcc()] + NVCC_FLAGS)


def _flags_changed(obj: str) -> bool:
    try:
        with open(obj + ".flags") as f:
            return f.read() != _flag_stamp()
    except OSError:
        return True


def build(force: bool = False, verbose: bool = False) -> str:
    srcs = sources()
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers += [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE)]
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        if force or _stale(obj, [src] + headers) or _flags_changed(obj):
            cmd = [_nvcc()] + NVCC_FLAGS + ["-c", src, "-o", obj]
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")
            if verbose:
                sys.stderr.write(res.stderr)
            with open(obj + ".log", "w") as f:
                f.write(res.stderr)
            with open(obj + ".flags", "w") as f:
                f.write(_flag_stamp())
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    if force or _stale(LIB, objs):
        cmd = [_nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
