"""CPU: the C-ABI library builds, loads and exports every symbol include/b2col.h declares; without a
GPU it refuses to create a context (there is no CPU path to fall back to)."""
import ctypes as C
import os
import re

import pytest
import torch

from conftest import ROOT
from engine_b200 import build as b2build
from engine_b200 import capi


@pytest.fixture(scope="module")
def lib():
    b2build.build()
    return capi.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b2col.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2col_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    names = declared_symbols()
    assert len(names) >= 20
    assert sorted(capi.SYMBOLS) == names, "capi.SYMBOLS must list exactly what include/b2col.h declares"
    for n in names:
        assert hasattr(lib, n), f"libb2col.so does not export {n}"
    assert lib.b2col_abi_version() == 4


def test_pod_sizes_match_reference():
    assert capi.RAY_DTYPE.itemsize == 24 and capi.RAYCAST_DTYPE.itemsize == 36
    assert capi.AABB_DTYPE.itemsize == 24 and capi.NODE_DTYPE.itemsize == 60


def test_sass_is_sm100a_without_fma_in_parity_kernels():
    """The library is compiled for sm_100a only, and the parity-critical kernels contain no fused
    multiply-add (the reference build has none)."""
    import subprocess

    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-sass", capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out and "sm_90" not in out
    cur = None
    ffma = {}
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
        elif cur and re.search(r"\bFFMA\b", line):
            ffma[cur] = ffma.get(cur, 0) + 1
    # The refit and transform kernels must contain no FFMA: there it could only be a contracted
    # multiply-add, which the reference build does not have.  k_trace_reftopo may contain FFMA only inside
    # the correctly-rounded IEEE division / square-root sequences of the ray prologue (load_ray: 1/sqrt(dot),
    # 3x 1/d expand to Newton steps) — a fixed, small count.  k_trace_lbvh additionally uses explicit fmaf in
    # its conservative 15-bit grid filter, which only prunes; its exact leaf tests go through
    # __fmul_rn/__fadd_rn/__fsub_rn (b2col_math.cuh), which ptxas never contracts.
    # k_refit_keys ends with the frame's header arithmetic (frame_tail -> finish_header -> make_grid: three IEEE divisions and
    # three reciprocals, whose Newton steps are FFMAs).  k_header is exactly that arithmetic and nothing else, so every
    # instantiation of k_refit_keys must contain precisely as many FFMAs as k_header: one contracted multiply-add in the
    # transform / refit / key arithmetic would show up as a surplus.
    header = [v for k, v in ffma.items() if "k_header" in k]
    assert len(header) == 1, ffma
    fused = {k: v for k, v in ffma.items() if "k_refit_keys" in k}
    assert len(fused) == 6 and all(v == header[0] for v in fused.values()), f"k_refit_keys holds FFMAs beyond the header arithmetic ({header[0]}): {fused}"
    bad = {k: v for k, v in ffma.items() if any(t in k for t in ("k_refit", "k_propagate")) and "k_refit_keys" not in k}
    assert not bad, f"FFMA found in parity-critical kernels: {bad}"
    ref = [v for k, v in ffma.items() if "k_trace_reftopo" in k]
    assert len(ref) == 1 and ref[0] <= 48, ffma


@pytest.mark.skipif(torch.cuda.is_available(), reason="GPU present")
def test_no_cpu_fallback(lib):
    h = C.c_void_p()
    st = lib.b2col_create(C.byref(h), None)
    assert st == capi.ERR_NO_DEVICE and not h.value
    assert b"no CPU path" in lib.b2col_last_error(None)
    from engine_b200.collision import CollisionSystem

    with pytest.raises(capi.B2ColError):
        CollisionSystem()


def test_expand_hits_restores_reference_raycasts(lib):
    """b2col_expand_hits (a pure host function: runs without a GPU) turns compact hits (entity, distance, hit | side << 8) back into
    the reference's Raycast: location = normalise(d) * t + o, normal from the side — against the golden vectors the real reference produced."""
    import numpy as np

    from conftest import assert_raycast_equal, load_golden

    g = load_golden("city2000")
    rays, want = np.ascontiguousarray(g["rays"], np.float32), g["raycast"]
    hits = np.zeros(rays.shape[0], capi.HIT_DTYPE)
    hits["hit_entity"] = want["hit_entity"]
    hits["distance"] = want["distance"]
    n = want["normal"]
    # engine::AABBSide { Left, Right, Bottom, Top, Front, Back } = -x, +x, -z, +z, -y, +y (reference src/system_collisions.cpp:175-194)
    side = np.select([n[:, 0] < 0, n[:, 0] > 0, n[:, 2] < 0, n[:, 2] > 0, n[:, 1] < 0, n[:, 1] > 0], [0, 1, 2, 3, 4, 5], 0).astype(np.uint32)
    hits["flags"] = want["hit"].astype(np.uint32) | (side << 8)
    for threads in (1, 4):
        out = np.empty(rays.shape[0], capi.RAYCAST_DTYPE)
        assert lib.b2col_expand_hits(capi.ptr(rays), capi.ptr(hits), rays.shape[0], capi.ptr(out), threads) == capi.OK
        assert_raycast_equal(out, want, f"expand_hits, {threads} threads")
    assert want["hit"].sum() > 100
