#!/usr/bin/env python
"""Dev tool: device-resident trace throughput of one workload (use B2COL_LIB=variants/... to compare kernel variants)."""
import argparse, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="config3", choices=["config3", "config5", "config2", "city", "hier"])
    ap.add_argument("--colliders", type=int, default=1_000_000)
    ap.add_argument("--rays", type=int, default=1 << 24)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--sort-rays", type=int, default=0, help="experiment: order the rays by (origin cell Morton code at this many bits per axis, direction octant) first")
    ap.add_argument("--aimed", action="store_true", help="rays aimed at colliders from ~50 units away (hit fraction > 99 %) instead of the workload's own rays")
    a = ap.parse_args()
    import torch
    from engine_b200 import synth
    from engine_b200.collision import CollisionSystem
    if a.workload == "config3":
        s, rays = synth.uniform_boxes(a.colliders), synth.random_rays(a.rays)
    elif a.workload == "config5":
        s, rays = synth.clustered_boxes(a.colliders), synth.random_rays(a.rays, seed=4, extent=2000.0)
    elif a.workload == "hier":
        s, rays = synth.hierarchy_scene(a.colliders, 1000), synth.random_rays(a.rays)
    else:
        s = synth.static_city(a.colliders)
        w = int(np.sqrt(a.rays))
        rays = synth.camera_rays(w, w) if a.workload == "config2" else synth.random_rays(a.rays, seed=4, extent=500.0)
    if a.aimed:
        rng = np.random.default_rng(58)
        target = s.pos[1 + rng.integers(0, s.n_colliders, a.rays)] + rng.uniform(-0.1, 0.1, (a.rays, 3)).astype(np.float32)
        origin = target + rng.standard_normal((a.rays, 3)).astype(np.float32) * 50.0
        rays = np.concatenate([origin, target - origin], axis=1).astype(np.float32)
    R = rays.shape[0]
    if a.sort_rays:
        b = a.sort_rays
        o = rays[:, 0:3]
        q = np.clip(((o - o.min(0)) / (o.max(0) - o.min(0)) * (1 << b)).astype(np.int64), 0, (1 << b) - 1)
        key = np.zeros(R, np.int64)
        for bit in range(b):
            for k in range(3):
                key |= ((q[:, k] >> bit) & 1) << (3 * bit + (2 - k))
        octant = (rays[:, 3] > 0).astype(np.int64) | ((rays[:, 4] > 0).astype(np.int64) << 1) | ((rays[:, 5] > 0).astype(np.int64) << 2)
        rays = np.ascontiguousarray(rays[np.argsort((key << 3) | octant, kind="stable")])
    cs = CollisionSystem(device=0)
    cs.set_scene(s)
    cs.on_update()
    d_rays = torch.from_numpy(rays).cuda()
    d_out = torch.empty(R * 36, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    for _ in range(3):
        cs.raycast_device(d_rays, R, d_out)
    cs.sync()
    cs.event_record(2)
    for _ in range(a.steps):
        cs.raycast_device(d_rays, R, d_out)
    cs.event_record(3)
    cs.sync()
    ms = cs.event_elapsed_ms(2, 3) / a.steps
    out = d_out.view(R, 36)
    print(json.dumps({"sort_rays": a.sort_rays, "lib": os.environ.get("B2COL_LIB", "default"), "workload": a.workload + (" aimed" if a.aimed else ""), "key_bits": cs.stats()["key_bits"], "colliders": s.n_colliders, "rays": R, "ms": ms, "mrays_s": R / ms / 1e3,
                      "hit_fraction": float((out[:, 32] != 0).float().mean().item()), "checksum": int(out[:, 24:28].contiguous().view(torch.int32).to(torch.int64).sum().item())}))

if __name__ == "__main__":
    main()
