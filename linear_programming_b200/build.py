"""In-tree build of liblp_b200.so (the C-ABI CUDA library) with nvcc for sm_100a.

    python -m linear_programming_b200.build [--force] [-v]

The .so is written next to this file (git-ignored; it travels to the GPU box with the
working tree).  nvcc cross-compiles without a GPU.  Every source is compiled to its own
object (in parallel, only when it or a header changed) and the objects are linked once.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJDIR = os.path.join(CSRC, "_obj")
LIB_PATH = os.path.join(HERE, "liblp_b200.so")
SOURCES = ["lp_api.cu", "simplex_batched.cu", "simplex_batched_variants.cu", "simplex_single.cu",
           "gauss_jordan_batched.cu", "verify_bounds.cu", "host_narrow.cpp"]
HEADERS = ["lp_common.cuh", "lp_host.h", "host_narrow.h", "simplex_batched_impl.cuh", "simplex_batched_variants.h",
           os.path.join("..", "..", "include", "lp_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    # no -use_fast_math, no -fmad tricks: IEEE division, and the data path uses
    # __dmul_rn/__dadd_rn explicitly so no product+sum is ever contracted into an FMA
    "--fmad=false",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build liblp_b200.so")


def _sources():
    return [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _header_mtime() -> float:
    deps = [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    return max(os.path.getmtime(d) for d in deps if os.path.exists(d))


def _obj(src: str) -> str:
    return os.path.join(OBJDIR, os.path.splitext(src)[0] + ".o")


def _obj_stale(src: str, hdr_t: float) -> bool:
    o = _obj(src)
    if not os.path.exists(o):
        return True
    t = os.path.getmtime(o)
    return os.path.getmtime(os.path.join(CSRC, src)) > t or hdr_t > t


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return _header_mtime() > t or any(os.path.getmtime(os.path.join(CSRC, s)) > t for s in _sources())


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB_PATH
    nvcc = find_nvcc()
    os.makedirs(OBJDIR, exist_ok=True)
    hdr_t = _header_mtime()
    todo = [s for s in _sources() if force or _obj_stale(s, hdr_t)]

    def compile_one(src: str):
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", "-o", _obj(src), os.path.join(CSRC, src)]
        return src, subprocess.run(cmd, capture_output=True, text=True)

    with ThreadPoolExecutor(max_workers=max(1, min(len(todo), os.cpu_count() or 1))) as pool:
        for src, res in pool.map(compile_one, todo):
            if res.returncode != 0:
                raise RuntimeError("nvcc failed on %s:\n%s%s" % (src, res.stdout, res.stderr))
            if verbose:
                sys.stderr.write(res.stderr)
    link = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC",
            "-o", LIB_PATH] + [_obj(s) for s in _sources()]
    res = subprocess.run(link, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + res.stdout + res.stderr)
    return LIB_PATH


def build_variant(name: str, defines, verbose: bool = False) -> str:
    """An experiment library next to the product one: variants/liblp_b200_<name>.so built with extra -D flags
    (A/B on one box through LPB_LIB_PATH, see tools/gpu/ab_variants.sh).  Always rebuilt."""
    nvcc = find_nvcc()
    vdir = os.path.join(HERE, "variants")
    odir = os.path.join(vdir, "_obj_" + name)
    os.makedirs(odir, exist_ok=True)
    out = os.path.join(vdir, "liblp_b200_%s.so" % name)

    def compile_one(src: str):
        obj = os.path.join(odir, os.path.splitext(src)[0] + ".o")
        cmd = [nvcc] + NVCC_FLAGS + list(defines) + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", "-o", obj, os.path.join(CSRC, src)]
        return src, obj, subprocess.run(cmd, capture_output=True, text=True)

    objs = []
    with ThreadPoolExecutor(max_workers=os.cpu_count() or 1) as pool:
        for src, obj, res in pool.map(compile_one, _sources()):
            if res.returncode != 0:
                raise RuntimeError("nvcc failed on %s:\n%s%s" % (src, res.stdout, res.stderr))
            if verbose:
                sys.stderr.write(res.stderr)
            objs.append(obj)
    res = subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC",
                          "-o", out] + objs, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + res.stdout + res.stderr)
    return out


if __name__ == "__main__":
    if "--variant" in sys.argv:  # python -m linear_programming_b200.build --variant timing -DLPB_TIMING
        i = sys.argv.index("--variant")
        print(build_variant(sys.argv[i + 1], [a for a in sys.argv[i + 2:] if a.startswith("-D")], verbose="-v" in sys.argv))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
