"""Builds augcl_b200/libaugcl_b200.so in-tree with nvcc for sm_100a (no torch, no JIT cache).

    python -m augcl_b200.build [--force]
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = [os.path.join(HERE, "csrc", f) for f in ("state.cu", "ring.cu", "fused.cu", "augment.cu", "augment2.cu", "rng.cu")]
HDRS = [os.path.join(HERE, "csrc", h) for h in ("common.cuh", "ring.cuh", "pixel.cuh")] + [os.path.join(ROOT, "include", "augcl_b200.h")]
LIB = os.path.join(HERE, "libaugcl_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "--fmad=false",            # parity: the reference's elementwise chains round every op
    "-Xcompiler", "-fPIC", "-shared", "-cudart", "static",
    "-I", os.path.join(ROOT, "include"),
]


def up_to_date() -> bool:
    if not os.path.isfile(LIB):
        return False
    t = os.path.getmtime(LIB)
    return all(os.path.getmtime(p) <= t for p in SRC + HDRS + [os.path.abspath(__file__)])


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and up_to_date():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + SRC
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
