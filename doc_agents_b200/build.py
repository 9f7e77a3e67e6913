"""Builds libdocagents_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m doc_agents_b200.build [--force] [--verbose]
"""
from __future__ import annotations

import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdocagents_b200.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC,-O2,-Wall,-Wno-unused-function", "-ccbin", "/usr/bin/g++",
    "--shared", "-cudart", "static",
]


def sources() -> list[str]:
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _deps() -> list[str]:
    return sources() + glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(INCLUDE, "*.h")) + [os.path.abspath(__file__)]


OBJ_DIR = os.path.join(os.path.dirname(HERE), "build", "obj")


def _headers() -> list[str]:
    return glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(INCLUDE, "*.h")) + [os.path.abspath(__file__)]


def build(force: bool = False, verbose: bool = False) -> str:
    """One object per .cu (compiled in parallel, only the stale ones), then one link.  Objects live under build/obj
    (git-ignored); the library is written aside and renamed, so concurrent builders never see a torn file."""
    if not force and os.path.exists(LIB) and all(os.path.getmtime(d) <= os.path.getmtime(LIB) for d in _deps()):
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    nvcc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "nvcc")
    extra = os.environ.get("DA_NVCC_EXTRA", "").split()   # extra nvcc flags for experiments (-D..., -Xptxas=...)
    os.makedirs(OBJ_DIR, exist_ok=True)
    tag = os.path.join(OBJ_DIR, ".flags")
    flags_now = " ".join(NVCC_FLAGS + extra)
    if not os.path.exists(tag) or open(tag).read() != flags_now:
        force = True
    hdr_time = max(os.path.getmtime(h) for h in _headers())
    compile_flags = [f for f in NVCC_FLAGS if f not in ("--shared",)]

    def compile_one(src: str) -> str:
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(src), hdr_time):
            return obj
        tmp_o = f"{obj}.{os.getpid()}.tmp"
        cmd = [nvcc, *compile_flags, *extra, "-I", INCLUDE, "-c", src, "-o", tmp_o]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), file=sys.stderr)
        try:
            subprocess.run(cmd, check=True)
            os.replace(tmp_o, obj)
        finally:
            if os.path.exists(tmp_o):
                os.remove(tmp_o)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    open(tag, "w").write(flags_now)
    tmp = f"{LIB}.{os.getpid()}.tmp"
    try:
        subprocess.run([nvcc, "--shared", "-cudart", "static", "-ccbin", "/usr/bin/g++",
                        "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp, *objs, "-lnccl"], check=True)
        os.replace(tmp, LIB)
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
