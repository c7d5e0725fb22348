"""Extra 1-GPU records of bench.py's JSON line: the honest dynamic and clustered numbers.

  config4   BASELINE configs[3] as SURVEY.md 8(d) defines it: 1M colliders under 1000 moving roots (depth 2, 10 % depth 3),
            the root TRS perturbed ON THE DEVICE every frame, TRS hand-over + b2col_update timed per frame with CUDA
            events, 100 frames, median and p99; then 4M rays (half camera, half random).
  config5   BASELINE configs[4] at its largest 1-GPU size (10M clustered AABBs, 640 MB BVH > L2), 16M rays AIMED at
            colliders (hit fraction > 50 %) next to the uniform random rays (which mostly miss), each with its own roofline.
  update_e2e (tools/bench_update_e2e.py) the per-frame update through the drop-in shims inside the reference's Scene.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def _trace(cs, torch, rays, steps=5, warmup=3):
    R = rays.shape[0]
    d_rays = torch.from_numpy(rays).cuda()
    d_out = torch.empty(R * 36, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    for _ in range(warmup):
        cs.raycast_device(d_rays, R, d_out)
    cs.event_record(2)
    for _ in range(steps):
        cs.raycast_device(d_rays, R, d_out)
    cs.event_record(3)
    cs.sync()
    ms = cs.event_elapsed_ms(2, 3) / steps
    hits = int((d_out.view(R, 36)[:, 32] != 0).sum().item())
    del d_rays, d_out
    return {"rays": R, "ms_per_step": ms, "value": R / ms / 1e3, "unit": "Mrays/s", "hit_fraction": hits / R}


def config4(torch, frames=100):
    from engine_b200 import synth
    from engine_b200.collision import CollisionSystem

    n, roots = 1_000_000, 1000
    s = synth.hierarchy_scene(n, roots)
    cs = CollisionSystem(use_cuda_graph=True)
    cs.set_scene(s)
    cs.on_update()
    ext = torch.cuda.ExternalStream(cs.stream)
    d_pos, d_rot, d_scl = (torch.from_numpy(a).cuda() for a in (s.pos, s.rot, s.scl))
    g = torch.Generator(device="cuda").manual_seed(7)
    dpos = 0.05 * torch.randn((8, roots, 3), device="cuda", generator=g)
    q = torch.randn((8, roots, 4), device="cuda", generator=g)
    q = q / q.norm(dim=2, keepdim=True)  # seeded re-rotation of the roots, as synth.perturb_roots
    torch.cuda.synchronize()
    out = {}
    with torch.cuda.stream(ext):
        for mode in ("rebuild", "refit_only", "rebuild_bound"):
            evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(frames + 5)]
            if mode == "rebuild_bound":
                cs.bind_trs_device(d_pos, d_rot, d_scl)  # zero-copy: the update reads these tensors in place; [0] -> [1] below is the animation alone
            for f in range(frames + 5):
                evs[f][0].record(ext)
                d_pos[1 : roots + 1] += dpos[f % 8]
                d_rot[1 : roots + 1] = q[f % 8]
                if mode != "rebuild_bound":
                    cs.set_trs_device(d_pos, d_rot, d_scl)  # 44 B/entity device-to-device into the context's SoA
                evs[f][1].record(ext)
                (cs.on_update_refit if mode == "refit_only" else cs.on_update)(wait=False)
                evs[f][2].record(ext)
            cs.sync()
            t = np.array([[e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])] for e in evs[5:]])
            out[mode] = {"frame_ms_median": float(np.median(t.sum(1))), "frame_ms_p99": float(np.percentile(t.sum(1), 99)),
                         "trs_handover_ms_median": float(np.median(t[:, 0])), "update_ms_median": float(np.median(t[:, 1])),
                         "update_ms_p99": float(np.percentile(t[:, 1], 99))}
    cs.set_trs_device(d_pos, d_rot, d_scl)  # back to the context's own copy
    cs.on_update()
    st = cs.stats()
    rays = np.concatenate([synth.camera_rays(2048, 1024, eye=(0.0, 0.0, 300.0), look=(1.0, 0.2, -0.3)), synth.random_rays(2 << 20, seed=44)])
    peak = _peak()
    ms = out["rebuild"]["frame_ms_median"]
    rec = {"workload": "configs[3]: 1M dynamic colliders under 1000 moving roots (depth 2, 10 % depth 3); root TRS perturbed on the device every frame; frame = TRS hand-over "
                       "(set_trs_device) + b2col_update (propagate, refit, full LBVH rebuild), per-frame CUDA events",
           "frames": frames, "entities": st["n_entities"], "colliders": st["n_colliders"], "levels": st["n_levels"], "sort_passes": st["sort_passes"],
           "rebuild": out["rebuild"], "refit_only": out["refit_only"],
           "rebuild_bound_trs": dict(out["rebuild_bound"], what="the same frames with the caller's TRS tensors bound (b2col_bind_trs_device: read in place, no hand-over copy); "
                                                                 "trs_handover_ms = the root animation alone"),
           "roofline": {"bound": "hbm", "algorithmic_bytes": 280.0 * n, "achieved": 280.0 * n / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": 280.0 * n / (ms * 1e-3) / 1e9 / peak},
           "rays_4m": _trace(cs, torch, rays)}
    cs.close()
    return rec


def config5(torch, n=10_000_000, n_rays=16 << 20):
    from engine_b200 import synth
    from engine_b200.collision import CollisionSystem

    s = synth.clustered_boxes(n, seed=5)
    cs = CollisionSystem(use_cuda_graph=True)
    cs.set_scene(s)
    cs.on_update()
    for _ in range(3):
        cs.on_update(wait=False)
    cs.event_record(0)
    for _ in range(10):
        cs.on_update(wait=False)
    cs.event_record(1)
    cs.sync()
    upd = cs.event_elapsed_ms(0, 1) / 10
    rng = np.random.default_rng(58)
    target = s.pos[1 + rng.integers(0, n, n_rays)] + rng.uniform(-0.1, 0.1, (n_rays, 3)).astype(np.float32)
    origin = target + rng.standard_normal((n_rays, 3)).astype(np.float32) * 50.0
    aimed = np.concatenate([origin, target - origin], axis=1).astype(np.float32)  # unnormalised directions: GetRaycast normalises
    del target, origin
    peak = _peak()
    blob = cs.bvh_blob()[1]
    rec = {"workload": f"configs[4] at its largest 1-GPU size: {n} clustered AABBs (1000 Gaussian clusters, Zipf populations, 5 % background, seed 5), {n_rays} rays; "
                       f"the {blob} byte BVH exceeds the 126 MB L2",
           "colliders": n, "bvh_bytes": blob, "update_ms": upd, "update_frac": 280.0 * n / (upd * 1e-3) / 1e9 / peak}
    # dram__bytes of these two launches from the committed ncu capture of THIS build of the trace kernel (tools/capture_traffic.py --config5)
    captured = {}
    try:
        import bench

        tr = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic_config5.json")))
        if tr.get("source_hash") == bench.source_hash() and (tr["colliders"], tr["rays"]) == (n, n_rays):
            captured = tr
    except (OSError, KeyError, ValueError, ImportError):
        pass
    for name, rays in (("aimed", aimed), ("random", None)):
        if rays is None:
            rays = synth.random_rays(n_rays, seed=4, extent=2000.0)
        r = _trace(cs, torch, rays, steps=3)
        algo = 60.0 * n_rays + 64.0 * (n - 1)
        r["roofline"] = {"bound": "hbm", "algorithmic_bytes": algo, "achieved": algo / (r["ms_per_step"] * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": algo / (r["ms_per_step"] * 1e-3) / 1e9 / peak, "traffic": None,
                         "note": "60 B/ray + the node set once"}
        if name in captured:
            t = captured[name]["dram_bytes_read"] + captured[name]["dram_bytes_write"]
            r["roofline"].update({"traffic": t, "traffic_gbs": t / (r["ms_per_step"] * 1e-3) / 1e9, "traffic_frac_of_peak": t / (r["ms_per_step"] * 1e-3) / 1e9 / peak,
                                  "l2_hit_rate_pct": captured[name].get("lts_sector_hit_rate_pct"),
                                  "traffic_source": f"ncu --set full capture of this build ({captured['source_hash']}), profiles/r2_traffic_config5.json"})
        rec[name] = r
        del rays
    cs.close()
    return rec


def run_all(cs_main, args, torch) -> dict:
    out = {}
    for name, fn in (("config4", config4), ("config5", config5)):
        try:
            out[name] = fn(torch)
        except Exception as e:  # an extra record must never cost the headline line
            out[name] = {"error": f"{type(e).__name__}: {e}"}
        torch.cuda.empty_cache()
    try:
        from tools import bench_update_e2e

        out["update_e2e"] = bench_update_e2e.run()
    except Exception as e:
        out["update_e2e"] = {"error": f"{type(e).__name__}: {e}"}
    return out


if __name__ == "__main__":
    import torch

    which = sys.argv[1:] or ["config4", "config5"]
    for w in which:
        print(json.dumps({w: globals()[w](torch)}), flush=True)
