"""b2col_raycast_batch_static — the entry point that allocates nothing (its device scratch is a file-scope static __device__
array; include/b2col.h).  No counterpart in the reference.

Collected LAST on purpose (test_zz_*): a failure here must not hide the tests before it.
The test calls the new entry point through the C ABI (ctypes) and compares its output with the existing entry point's
(b2col_raycast_batch) under the bar every GPU test holds (conftest.assert_raycast_equal: flags and entity IDs bit-exact,
floats numerically equal), on several calls in a row — the static array keeps the previous call's contents."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conftest import assert_raycast_equal  # noqa: E402

pytestmark = pytest.mark.gpu

STATIC_MAX_RAYS = 65536  # B2COL_STATIC_MAX_RAYS, include/b2col.h


def _static(cs, rays):
    from engine_b200 import capi

    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    out = np.full(rays.shape[0], 0xAB, np.uint8).repeat(36).view(capi.RAYCAST_DTYPE)  # poisoned: every byte must be written by the call
    assert out.shape[0] == rays.shape[0]
    st = cs.lib.b2col_raycast_batch_static(cs.h, capi.ptr(rays), rays.shape[0], capi.ptr(out))
    return st, out


def test_static_scratch_entry_point_matches_the_allocating_one():
    import torch

    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    from engine_b200 import capi, synth
    from engine_b200.collision import CollisionSystem

    cs = CollisionSystem()
    # before the first update every ray misses, as through b2col_raycast_batch
    rays0 = synth.random_rays(1000, seed=11, extent=300)
    st, got = _static(cs, rays0)
    assert st == capi.OK and not got.view(np.uint8).any()

    s = synth.hierarchy_scene(40_000, 60, seed=21, extent=300)
    cs.set_scene(s)
    cs.on_update()
    # call 1: a full-size batch; call 2: a smaller one (the array still holds call 1's rays and results behind it);
    # call 3: the maximum size again with other rays; call 4: a single ray
    batches = [np.concatenate([synth.random_rays(50_000, seed=31, extent=320), synth.axis_rays(5_000, seed=32, extent=300)]),
               synth.random_rays(7_777, seed=33, extent=320),
               synth.random_rays(STATIC_MAX_RAYS, seed=34, extent=320),
               synth.random_rays(1, seed=35, extent=50)]
    for k, rays in enumerate(batches):
        want = cs.get_raycasts(rays)  # b2col_raycast_batch: the existing entry point
        st, got = _static(cs, rays)
        assert st == capi.OK, cs.lib.b2col_last_error(cs.h)
        assert_raycast_equal(got, want, f"call {k + 1} ({rays.shape[0]} rays): static-scratch entry point vs b2col_raycast_batch")
        assert (got.view(np.uint8) == want.view(np.uint8)).all() or (got["distance"] == want["distance"]).all()
    assert want.shape[0] == 1 and sum(int(cs.get_raycasts(b)["hit"].sum()) for b in batches[:3]) > 1000, "the batches should hit something"

    # a moved scene, rebuilt (adaptive key width / plan-before-frame builds included): still the same answers
    for frame in range(1, 4):
        synth.perturb_roots(s, 60, frame)
        cs.set_trs(s.pos, s.rot, s.scl)
        cs.on_update()
        st, got = _static(cs, batches[0])
        assert st == capi.OK
        assert_raycast_equal(got, cs.get_raycasts(batches[0]), f"frame {frame}: static-scratch entry point vs b2col_raycast_batch")

    # a second context on the same GPU shares the array (sequential calls are fine)
    cs2 = CollisionSystem()
    cs2.set_scene(synth.uniform_boxes(5_000, extent=100.0))
    cs2.on_update()
    r2 = synth.random_rays(20_000, seed=36, extent=110)
    st, got2 = _static(cs2, r2)
    assert st == capi.OK
    assert_raycast_equal(got2, cs2.get_raycasts(r2), "second context")
    st, got = _static(cs, batches[1])
    assert st == capi.OK
    assert_raycast_equal(got, cs.get_raycasts(batches[1]), "first context again")

    # beyond the maximum: a clear error, nothing written
    too_many = synth.random_rays(STATIC_MAX_RAYS + 1, seed=37, extent=320)
    st, out = _static(cs, too_many)
    assert st == capi.ERR_INVALID
    msg = cs.lib.b2col_last_error(cs.h).decode()
    assert "B2COL_STATIC_MAX_RAYS" in msg and str(STATIC_MAX_RAYS) in msg, msg
    assert (out.view(np.uint8) == 0xAB).all()
    # n = 0 and null arguments
    assert cs.lib.b2col_raycast_batch_static(cs.h, None, 0, None) == capi.OK
    assert cs.lib.b2col_raycast_batch_static(cs.h, None, 5, None) == capi.ERR_INVALID
    cs2.close()
    cs.close()
