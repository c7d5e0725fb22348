import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    from oracle.oracle import Port

    return Port()


def load_golden(name):
    from engine_b200.synth import Scene
    from oracle.oracle import NODE_DTYPE, RAYCAST_DTYPE

    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    scene = Scene(parent=z["parent"], pos=z["pos"], rot=z["rot"], scl=z["scl"], collider_entity=z["collider_entity"],
                  collider_aabb=z["collider_aabb"])
    return dict(
        scene=scene,
        world=z["world"],
        world_aabb_entity=z["world_aabb_entity"],
        world_aabb=z["world_aabb"],
        bvh=np.ascontiguousarray(z["bvh"]).view(NODE_DTYPE).reshape(-1),
        rays=z["rays"],
        raycast=np.ascontiguousarray(z["raycast"]).view(RAYCAST_DTYPE).reshape(-1),
    )


SCENE_FIXTURES = ["game600", "hier1500", "grid512", "city2000", "single", "pair", "triple"]


def assert_raycast_equal(got, want, what="", ignore_entity_on=None, rel_tol=0.0, nan_equal=False):
    """Parity bar (BASELINE.json north_star): hit flag and hit_entity bit-exact; distance, location,
    normal within 1e-5 relative.  We hold the floats to the stricter numeric equality by default
    (rel_tol=0): == treats +0 and -0 as equal, the one thing that legitimately differs — the
    reference's fminf/fmaxf return an operand-order-dependent zero sign (libm call with
    compiler-chosen argument order), so the sign of a zero distance is not defined by its source.
    ignore_entity_on: boolean mask of rays (exact tmin ties) whose hit_entity may differ.
    nan_equal: degenerate inputs only — a NaN matches a NaN (a NaN ray origin gives a NaN location on both sides)."""
    assert got.shape == want.shape, (what, got.shape, want.shape)
    bad = np.nonzero(got["hit"] != want["hit"])[0]
    assert bad.size == 0, f"{what}: hit flag differs on {bad.size} rays, first {bad[:5]}"
    ent_bad = got["hit_entity"] != want["hit_entity"]
    if ignore_entity_on is not None:
        ent_bad &= ~ignore_entity_on
    bad = np.nonzero(ent_bad)[0]
    assert bad.size == 0, f"{what}: hit_entity differs on {bad.size} rays, first {bad[:5]}: {got[bad[:3]]} vs {want[bad[:3]]}"
    for f in ("distance", "location", "normal"):
        a, b = got[f].astype(np.float64), want[f].astype(np.float64)
        if rel_tol == 0.0:
            ok = a == b
        else:
            ok = np.abs(a - b) <= rel_tol * np.maximum(np.abs(a), np.abs(b))
        if nan_equal:
            ok |= np.isnan(a) & np.isnan(b)
        if ignore_entity_on is not None and f == "normal":
            ok |= ignore_entity_on.reshape((-1,) + (1,) * (ok.ndim - 1))
        assert ok.all(), f"{what}: {f} differs on {int((~ok).sum())} values, first ray {np.nonzero(~ok)[0][:5]}"
    assert (got["pad"] == 0).all(), f"{what}: padding bytes not zero"
