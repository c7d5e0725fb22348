"""TEST INFRASTRUCTURE — run the reference's OWN O(N^2) BuildNode once at BASELINE config-2 size and cache the tree.

    python -m oracle.build_ref_tree [--n 100000] [--out oracle/_ref/config2_ref_bvh.npz]

The unmodified reference (oracle/_ref/libref_engine.so: Scene::Update = TransformSystem::onUpdate +
CollisionSystem::onUpdate, reference src/system_collisions.cpp:112-158,205-376) builds the BVH of
engine_b200.synth.static_city(n) — about half an hour of one core at n = 100 000 (BASELINE.md section 2:
x4 per doubling).  The result (its public `bvh_`, world AABBs and a fingerprint of the input scene) goes
to the git-ignored oracle/_ref/ and travels to the GPU box with the snapshot; tests/test_gpu_parity.py
uploads it for the reference-topology mode at config-2 size and skips loudly when it is absent.
Needs /root/reference (build time only).
"""
from __future__ import annotations

import argparse
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

DEFAULT_OUT = os.path.join(ROOT, "oracle", "_ref", "config2_ref_bvh.npz")


def scene_fingerprint(scene) -> str:
    h = hashlib.sha256()
    for a in (scene.parent, scene.pos, scene.rot, scene.scl, scene.collider_entity, scene.collider_aabb):
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100_000)
    ap.add_argument("--out", default=DEFAULT_OUT)
    args = ap.parse_args()
    from engine_b200 import synth
    from oracle.oracle import RefEngine

    scene = synth.static_city(args.n)
    ref = RefEngine()
    ref.create_entities(scene.parent[1:], scene.pos[1:], scene.rot[1:], scene.scl[1:])
    ref.add_colliders(scene.collider_entity, scene.collider_aabb)
    t0 = time.time()
    secs = ref.scene_update()  # TransformSystem + refit + BuildNode, the reference's own frame
    bvh = ref.bvh()
    ent, box, _ = ref.refit(scene.n_colliders)
    tmp = args.out + ".tmp.npz"
    np.savez_compressed(tmp, bvh=bvh.view(np.uint8), world_aabb_entity=ent, world_aabb=box.view(np.float32).reshape(-1, 6), n=np.int64(args.n),
                        fingerprint=np.frombuffer(scene_fingerprint(scene).encode(), np.uint8), build_secs=np.float64(secs))
    os.replace(tmp, args.out)
    print(f"reference BuildNode at n={args.n}: {secs:.1f} s ({time.time() - t0:.1f} s wall), {bvh.shape[0]} nodes -> {args.out}")


if __name__ == "__main__":
    main()
