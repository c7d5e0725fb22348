"""Measure the BASELINE.json configs that bench.py's headline line does not cover (configs[0], [1], [3], [4] at their
1-GPU sizes), one JSON line each.  bench.py stays the contract; this is the supplementary table of DESIGN.md section 6.

    python tools/bench_configs.py [config1 config2 config4 config5]
"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from engine_b200 import synth  # noqa: E402
from engine_b200.collision import CollisionSystem  # noqa: E402
from oracle.oracle import DropinEngine, Port, RefEngine  # noqa: E402


def trace_rate(cs, rays, steps=5, warmup=3):
    """device-resident Mrays/s (CUDA events on the ctx stream) and e2e Mrays/s through pinned host buffers"""
    R = rays.shape[0]
    h_rays = torch.from_numpy(rays).pin_memory()
    h_out = torch.empty(R * 36, dtype=torch.uint8).pin_memory()
    d_rays, d_out = h_rays.cuda(), torch.empty(R * 36, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    for _ in range(warmup):
        cs.raycast_device(d_rays, R, d_out)
    cs.event_record(2)
    for _ in range(steps):
        cs.raycast_device(d_rays, R, d_out)
    cs.event_record(3)
    dev_ms = cs.event_elapsed_ms(2, 3) / steps
    cs.raycast_host_pinned(h_rays, R, h_out)
    t0 = time.perf_counter()
    for _ in range(steps):
        cs.raycast_host_pinned(h_rays, R, h_out)
    e2e_ms = (time.perf_counter() - t0) * 1e3 / steps
    hits = int((d_out.view(R, 36)[:, 32] != 0).sum().item())
    return {"rays": R, "device_Mrays_s": R / dev_ms / 1e3, "device_ms": dev_ms, "e2e_Mrays_s": R / e2e_ms / 1e3, "e2e_ms": e2e_ms, "hit_fraction": hits / R}


def update_times(cs, frames=50):
    for _ in range(5):
        cs.on_update(wait=False)
    cs.event_record(0)
    for _ in range(frames):
        cs.on_update(wait=False)
    cs.event_record(1)
    return cs.event_elapsed_ms(0, 1) / frames


def cpu_reference_rays(scene, rays, seconds=8.0):
    """reference GetRaycast on all host threads over a binned-SAH tree injected into bvh_ (its own builder is O(N^2))"""
    if not RefEngine.available():
        return None
    port, ref = Port(), RefEngine()
    ref.create_entities(scene.parent[1:], scene.pos[1:], scene.rot[1:], scene.scl[1:])
    ref.add_colliders(scene.collider_entity, scene.collider_aabb)
    ref.transform_update()
    ent, box, _ = ref.refit(scene.n_colliders)
    ref.set_bvh(port.build_binned_sah(ent, box.view(np.float32).reshape(-1, 6)))
    threads = os.cpu_count() or 1
    n = min(len(rays), 20000)
    _, secs = ref.raycast(rays[:n], nthreads=threads, want_secs=True)
    n = int(min(len(rays), max(n, n * seconds / max(secs, 1e-6))))
    _, secs = ref.raycast(rays[:n], nthreads=threads, want_secs=True)
    return {"Mrays_s": n / secs / 1e6, "threads": threads, "sample_rays": n}


def config1():
    """test game scene, ~1k colliders, the controller's 20+1 rays per frame (reference CPU; and the drop-in on the GPU)"""
    s = synth.game_scene(1000)
    inp = synth.game_inputs(120)
    out = {"config": "configs[0]: game scene, 1000 colliders, 20+1 controller rays per frame", "frames": len(inp)}
    if RefEngine.available():
        e = RefEngine()
        e.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
        e.add_colliders(s.collider_entity, s.collider_aabb)
        e.add_camera((8.0, 2.0, 4.0))
        t0 = time.perf_counter()
        e.run_game(inp[:12])  # the reference rebuilds its O(N^2) tree every frame: ~0.3 s per frame at 1k colliders
        out["reference_ms_per_frame"] = (time.perf_counter() - t0) * 1e3 / 12
    if DropinEngine.available():
        for batched in (False, True):
            e = DropinEngine()
            e.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
            e.add_colliders(s.collider_entity, s.collider_aabb)
            e.add_camera((8.0, 2.0, 4.0), batched=batched)
            e.run_game(inp[:10])
            t0 = time.perf_counter()
            e.run_game(inp)
            out["dropin_batched_ms_per_frame" if batched else "dropin_ms_per_frame"] = (time.perf_counter() - t0) * 1e3 / len(inp)
    return out


def config2():
    s = synth.static_city(100_000)
    rays = synth.camera_rays(1000, 1000)
    cs = CollisionSystem(use_cuda_graph=True)
    cs.set_scene(s)
    cs.on_update()
    out = {"config": "configs[1]: 100k static AABBs, 1000x1000 coherent camera rays, 1 B200", "update_ms": update_times(cs)}
    out.update(trace_rate(cs, rays))
    out["cpu_reference"] = cpu_reference_rays(s, rays)
    return out


def config4():
    n, roots, frames = 1_000_000, 1000, 100
    s = synth.hierarchy_scene(n, roots)
    cs = CollisionSystem(use_cuda_graph=True)
    cs.set_scene(s)
    cs.on_update()
    d_pos, d_rot, d_scl = (torch.from_numpy(a).cuda() for a in (s.pos, s.rot, s.scl))
    g = torch.Generator(device="cuda").manual_seed(7)
    torch.cuda.synchronize()
    res = {}
    for mode in ("rebuild", "refit_only"):
        t = []
        for f in range(frames + 5):
            d_pos[1 : roots + 1] += 0.05 * torch.randn((roots, 3), device="cuda", generator=g)  # the frame's animation: roots drift
            torch.cuda.synchronize()
            cs.event_record(0)
            cs.set_trs_device(d_pos, d_rot, d_scl)  # 44 MB device-to-device into the context's SoA
            cs.event_record(1)
            (cs.on_update if mode == "rebuild" else cs.on_update_refit)(wait=False)
            cs.event_record(2)
            cs.sync()
            if f >= 5:
                t.append((cs.event_elapsed_ms(0, 1), cs.event_elapsed_ms(1, 2)))
        t = np.array(t)
        res[mode] = {"set_trs_device_ms_median": float(np.median(t[:, 0])), "update_ms_median": float(np.median(t[:, 1])), "update_ms_p99": float(np.percentile(t[:, 1], 99)),
                     "update_ms_max": float(t[:, 1].max())}
    cs.on_update()
    rays = np.concatenate([synth.camera_rays(2048, 1024, eye=(0.0, 0.0, 300.0), look=(1.0, 0.2, -0.3)), synth.random_rays(2 << 20, seed=44)])
    out = {"config": "configs[3]: 1M dynamic colliders under 1000 moving roots (depth 3), per-frame propagate + refit + full rebuild, 4M rays (half camera, half random)",
           "frames": frames, "stats": {k: cs.stats()[k] for k in ("n_entities", "n_colliders", "n_levels", "sort_passes", "graph_replays")}}
    out.update(res)
    out.update(trace_rate(cs, rays))
    if RefEngine.available():  # the reference's own TransformSystem + refit loop at this size (its builder cannot run: O(N^2))
        ref = RefEngine()
        ref.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
        ref.add_colliders(s.collider_entity, s.collider_aabb)
        ref.transform_update()
        tt = min(ref.transform_update() for _ in range(3))
        _, _, tr = ref.refit(n)
        out["cpu_reference"] = {"transform_update_ms": tt * 1e3, "refit_loop_ms": tr * 1e3, "threads": 1, "build": "not run: BuildNode is O(N^2), est. 53 h at 1M (SURVEY.md 8c)"}
    return out


def config5():
    n = 10_000_000
    s = synth.clustered_boxes(n, seed=5)
    cs = CollisionSystem(use_cuda_graph=True)
    cs.set_scene(s)
    cs.on_update()
    out = {"config": "configs[4] at 1 GPU: 10M clustered AABBs (640 MB BVH > L2), 16M incoherent rays", "update_ms": update_times(cs, 20), "sort_passes": cs.stats()["sort_passes"]}
    out.update(trace_rate(cs, synth.random_rays(16 << 20, seed=4, extent=2000.0), steps=3))
    return out


if __name__ == "__main__":
    which = sys.argv[1:] or ["config1", "config2", "config4", "config5"]
    for w in which:
        print(json.dumps(globals()[w]()), flush=True)
