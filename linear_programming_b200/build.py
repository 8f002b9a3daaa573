"""In-tree build of liblp_b200.so (the C-ABI CUDA library) with nvcc for sm_100a.

    python -m linear_programming_b200.build [--force]

The .so is written next to this file (git-ignored; it travels to the GPU box with the
working tree).  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "liblp_b200.so")
SOURCES = ["lp_api.cu", "simplex_batched.cu", "simplex_single.cu", "gauss_jordan_batched.cu", "verify_bounds.cu", "host_narrow.cpp"]
HEADERS = ["lp_common.cuh", "lp_host.h", "host_narrow.h", "simplex_batched_impl.cuh", os.path.join("..", "..", "include", "lp_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
    # no -use_fast_math, no -fmad tricks: IEEE division, and the data path uses
    # __dmul_rn/__dadd_rn explicitly so no product+sum is ever contracted into an FMA
    "--fmad=false",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build liblp_b200.so")


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB_PATH
    cmd = [find_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB_PATH] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        sys.stderr.write(res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
