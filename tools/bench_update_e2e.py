"""update_e2e: the per-frame update THROUGH THE DROP-IN, host side included.

The reference's own Scene::Update loop runs with both shims compiled in (oracle/_ref/libdropin2_engine.so: engine_b200/shim's
TransformSystem + CollisionSystem inside the unmodified Scene / ECS) over E entities, 10 % of which move every frame (the rest
carry TransformComponent::is_static); next to it the UNMODIFIED reference's TransformSystem::onUpdate + refit loop on one host
thread (its BuildNode is O(N^2) and cannot run at these sizes: stated, not timed).  ms per frame, wall clock around
Scene::Update (the call returns with world_matrix written back; the GPU build is enqueued and overlaps the next frame's host
work, so one extra b2col_sync per frame is timed too: `frame_ms_synced`).
"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _scene(n, seed=11):
    from engine_b200 import synth

    s = synth.uniform_boxes(n, seed=seed, extent=1000.0 * (n / 1e6) ** (1 / 3))
    rng = np.random.default_rng(seed + 1)
    s.rot[1:] = synth.random_unit_quats(rng, n)
    moving = np.sort(rng.choice(n, n // 10, replace=False)).astype(np.uint32) + 1  # entity IDs
    return s, moving


def _frames(engine, s, moving, frames, rng, sync=None):
    t_frame = []
    pos = s.pos[moving].copy()
    for f in range(frames):
        pos += rng.uniform(-0.05, 0.05, pos.shape).astype(np.float32)
        engine.set_trs(moving, pos, s.rot[moving], s.scl[moving])  # the game moving its dynamic entities (not timed)
        t0 = time.perf_counter()
        engine.scene_update()
        t1 = time.perf_counter()
        if sync:
            sync()
        t_frame.append((t1 - t0, time.perf_counter() - t0))
    return np.array(t_frame) * 1e3


def one_size(n, frames=30):
    from oracle.oracle import Dropin2Engine, RefEngine

    s, moving = _scene(n)
    static = np.setdiff1d(np.arange(1, n + 1, dtype=np.uint32), moving)
    rec = {"entities": n, "colliders": n, "moving_per_frame": int(moving.shape[0]), "frames": frames}
    # ---- drop-in
    e = Dropin2Engine()
    t0 = time.perf_counter()
    e.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
    e.add_colliders(s.collider_entity, s.collider_aabb)
    e.set_static(static, True)
    rec["scene_setup_s"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    e.scene_update()  # first frame: everything is packed and uploaded, the tree is built
    rec["dropin_first_frame_ms"] = (time.perf_counter() - t0) * 1e3
    rng = np.random.default_rng(5)
    _frames(e, s, moving, 3, rng)
    for label, opts in (("dropin", {}), ("dropin_boxes_immutable", {"collider_check_interval": 0}), ("dropin_no_writeback", {"collider_check_interval": 0, "write_back_world_matrices": 0})):
        for k, v in {"collider_check_interval": 1, "write_back_world_matrices": 1, **opts}.items():
            e.shim_option(k, v)
        t = _frames(e, s, moving, frames, rng, sync=lambda: e.raycast_batch(np.array([[0, 0, 0, 1, 0, 0]], np.float32)))
        rec[label] = {"frame_ms_median": float(np.median(t[:, 0])), "frame_ms_p90": float(np.percentile(t[:, 0], 90)), "options": opts}
    rec["dropin"]["full_uploads"] = e.shim_option("full_uploads")
    rec["dropin"]["sparse_uploads"] = e.shim_option("sparse_uploads")
    # world matrices / hits identical to the reference on a sample frame are asserted by tests/test_gpu_parity.py (test_dropin_*)
    del e
    # ---- the unmodified reference, one thread (as it runs): TransformSystem::onUpdate + the refit loop; BuildNode not runnable
    if RefEngine.available():
        r = RefEngine()
        r.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
        r.add_colliders(s.collider_entity, s.collider_aabb)
        tt, tr = [], []
        for _ in range(3):
            tt.append(r.transform_update())
            tr.append(r.refit(n)[2])
        rec["reference"] = {"transform_update_ms": float(np.median(tt)) * 1e3, "refit_loop_ms": float(np.median(tr)) * 1e3, "threads": 1,
                            "frame_ms": (float(np.median(tt)) + float(np.median(tr))) * 1e3,
                            "build": "not run: BuildNode is O(N^2) (est. 32 min at 100k, 53 h at 1M, SURVEY.md 8c); the frame figure is a LOWER bound of the reference's frame"}
        rec["speedup_vs_reference_lower_bound"] = rec["reference"]["frame_ms"] / rec["dropin"]["frame_ms_median"]
        del r
    return rec


def run(sizes=(100_000, 1_000_000)):
    from oracle.oracle import Dropin2Engine

    if not Dropin2Engine.available():
        return {"unavailable": "oracle/_ref/libdropin2_engine.so not built (needs /root/reference at build time)"}
    out = {"what": "Scene::Update with engine_b200/shim's TransformSystem + CollisionSystem compiled into the reference's Scene (libdropin2_engine.so), 10 % of the entities "
                   "moving per frame, the rest is_static; wall clock around Scene::Update, ms per frame (median); reference = its own TransformSystem::onUpdate + refit loop, one thread",
           "unit": "ms/frame"}
    for n in sizes:
        out[str(n)] = one_size(n)
    return out


if __name__ == "__main__":
    sizes = tuple(int(a) for a in sys.argv[1:]) or (100_000, 1_000_000)
    print(json.dumps(run(sizes)))
