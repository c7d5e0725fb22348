"""Sweep the trace kernel's tuning knobs on one GPU (dev tool; numbers go to stdout)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from engine_b200 import synth  # noqa: E402
from engine_b200.collision import CollisionSystem  # noqa: E402


def main():
    n_col = int(os.environ.get("N_COL", 1_000_000))
    n_rays = int(os.environ.get("N_RAYS", 8 * 1024 * 1024))
    s = synth.uniform_boxes(n_col)
    rays = synth.random_rays(n_rays)
    base = CollisionSystem()
    base.set_scene(s)
    base.on_update()
    p, nbytes = base.bvh_blob()
    d_rays = torch.from_numpy(rays).cuda()
    d_out = torch.empty(n_rays * 36, dtype=torch.uint8, device="cuda")
    ref = None
    sweeps = [x for x in os.environ.get("SWEEP", "thr=20").split(";") if x]
    for sw in sweeps:
        for kv in sw.split(","):
            k, v = kv.split("=")
            os.environ[{"thr": "B2COL_TRACE_FETCH_THRESHOLD", "bps": "B2COL_TRACE_BLOCKS_PER_SM"}[k]] = v
        cs = CollisionSystem()
        cs.bvh_import_blob(p, nbytes)
        for _ in range(2):
            cs.raycast_device(d_rays, n_rays, d_out)
        cs.sync()
        cs.event_record(0)
        for _ in range(3):
            cs.raycast_device(d_rays, n_rays, d_out)
        cs.event_record(1)
        ms = cs.event_elapsed_ms(0, 1) / 3
        got = d_out.cpu().numpy()
        if ref is None:
            ref = got
        print(f"{sw:24s} {ms:8.3f} ms  {n_rays / ms / 1e3:8.1f} Mrays/s  same={bool((got == ref).all())}", flush=True)
        cs.close()


if __name__ == "__main__":
    main()
