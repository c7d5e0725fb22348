"""Build libb2col.so (the C-ABI CUDA library) in-tree for sm_100a with nvcc.

    python -m engine_b200.build [--force]

The .so is git-ignored but travels to the GPU box with the repo snapshot.  nvcc cross-compiles
without a GPU.  -fmad=false: the parity-critical arithmetic already uses the _rn intrinsics, the
flag additionally keeps ptxas from contracting anything else on this path into FMAs.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libb2col.so")
SOURCES = ["b2col.cu"]
HEADERS = ["b2col_math.cuh", "b2col_bvh.cuh", "b2col_sort.cuh", "b2col_build.inl", "b2col_trace.inl", "b2col_comm.inl", "../../include/b2col.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo", "-fmad=false",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared",  # host arithmetic (b2col_expand_hits) rounds like the device's: no FMA
    "--expt-relaxed-constexpr",
]


# host-only C++: collider ingestion from .glb files (include/b2col_gltf.h)
HOST = os.path.join(HERE, "host")
GLTF_LIB = os.path.join(HERE, "libb2col_gltf.so")
GLTF_SOURCES = ["gltf_colliders.cpp", "gltf_capi.cpp"]
GLTF_HEADERS = ["gltf_colliders.h", "../../include/b2col_gltf.h"]


def build_gltf(force: bool = False) -> str:
    if not force and os.path.exists(GLTF_LIB):
        t = os.path.getmtime(GLTF_LIB)
        if not any(os.path.getmtime(os.path.join(HOST, f)) > t for f in GLTF_SOURCES + GLTF_HEADERS):
            return GLTF_LIB
    cxx = os.environ.get("CXX", "g++")
    # -ffp-contract=off: DecomposeTransform's sums of squares must round like the reference build's (no FMA)
    subprocess.check_call([cxx, "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-Wall", "-Wextra", "-o", GLTF_LIB,
                           *[os.path.join(HOST, s) for s in GLTF_SOURCES]])
    return GLTF_LIB


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False, out: str | None = None) -> str:
    global LIB
    if out:
        LIB, force = out, True
    if not out:
        build_gltf(force)
    if not force and not stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("B2COL_DEFINES", "").split()  # dev: e.g. "-DB2_SMEM_STACK=16 -DB2_NODE_LOAD_128"
    cmd = [nvcc, *NVCC_FLAGS, *extra, "-o", LIB, *[os.path.join(CSRC, s) for s in SOURCES], "-lcudart"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    out = next((a.split("=", 1)[1] for a in sys.argv if a.startswith("--out=")), None)
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, out=out))
