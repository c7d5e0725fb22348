#!/usr/bin/env python
"""Dev experiment: does a compact (breadth-first) placement of the top levels of the tree speed up the traversal of
trees larger than L2?  Nodes live at their split position, so the few thousand hot nodes of the upper levels are
scattered over the whole node array (one 2 MB page and one L2 line each).  This script permutes node indices on
the host so that the top `--levels` levels occupy indices [0, M) in breadth-first order, re-imports the blob and
times the unchanged trace kernel.

  python tools/exp_top_layout.py --workload config5 --colliders 10000000 --levels 16
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LEAF_BIT = 0x80000000


def bfs_top(q, root, levels):
    """node indices of the top `levels` levels in breadth-first order"""
    order, frontier = [], np.array([root], np.int64)
    for _ in range(levels):
        if frontier.size == 0:
            break
        order.append(frontier)
        kids = q[frontier, 6:8].reshape(-1)
        frontier = kids[(kids & LEAF_BIT) == 0].astype(np.int64)
    return np.concatenate(order)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="config5", choices=["config3", "config5"])
    ap.add_argument("--colliders", type=int, default=10_000_000)
    ap.add_argument("--rays", type=int, default=1 << 24)
    ap.add_argument("--levels", type=int, nargs="+", default=[12, 16, 18])
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    import torch

    from engine_b200 import synth
    from engine_b200.collision import CollisionSystem

    if a.workload == "config3":
        s, rays = synth.uniform_boxes(a.colliders), synth.random_rays(a.rays)
    else:
        s, rays = synth.clustered_boxes(a.colliders), synth.random_rays(a.rays, seed=4, extent=2000.0)
    R = rays.shape[0]
    os.environ["B2COL_TOP_LEVELS"] = "0"  # the production compact copy off: this script measures the layout itself
    cs = CollisionSystem(device=0)
    cs.set_scene(s)
    cs.on_update()
    d_rays = torch.from_numpy(rays).cuda()
    d_out = torch.empty(R * 36, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()

    def time_trace(c):
        for _ in range(3):
            c.raycast_device(d_rays, R, d_out)
        c.sync()
        c.event_record(2)
        for _ in range(a.steps):
            c.raycast_device(d_rays, R, d_out)
        c.event_record(3)
        c.sync()
        return c.event_elapsed_ms(2, 3) / a.steps

    base_ms = time_trace(cs)
    want = d_out.clone()
    _, nbytes = cs.bvh_blob()
    blob = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    cs.bvh_export_blob(blob, nbytes)
    cs.sync()
    words = blob.cpu().numpy().view(np.uint32).copy()
    n = int(words[1])
    q = words[64 : 64 + 8 * (n - 1)].reshape(n - 1, 8)
    root = int(words[3])
    res = {"lib": os.environ.get("B2COL_LIB", "default"), "workload": a.workload, "colliders": n, "rays": R, "node_order": "split position (production)", "ms": base_ms,
           "mrays_s": R / base_ms / 1e3}
    print(json.dumps(res), flush=True)
    for levels in a.levels:
        top = bfs_top(q, root, levels)
        M = top.shape[0]
        # permutation: top[i] -> i; the nodes displaced from [0, M) take the vacated indices
        perm = np.arange(n - 1, dtype=np.int64)
        in_low = np.zeros(n - 1, bool)
        in_low[top[top < M]] = True
        displaced = np.nonzero(~in_low[:M])[0]  # indices < M that are not top nodes
        vacated = top[top >= M]                  # top nodes living above M
        perm[top] = np.arange(M)
        perm[displaced] = vacated
        q2 = np.empty_like(q)
        q2[perm] = q
        for w in (6, 7):
            r = q2[:, w]
            internal = (r & LEAF_BIT) == 0
            r[internal] = perm[r[internal]].astype(np.uint32)
        out = words.copy()
        out[3] = perm[root]
        out[25], out[26] = perm[root], 0  # trace_root, top_count: this experiment moves the nodes themselves
        out[64 : 64 + 8 * (n - 1)] = q2.reshape(-1)
        blob2 = torch.from_numpy(out.view(np.uint8)).cuda()
        torch.cuda.synchronize()
        cs2 = CollisionSystem(device=0)
        cs2.bvh_import_blob(blob2, nbytes)
        ms = time_trace(cs2)
        same = bool((d_out == want).all().item())
        print(json.dumps({**res, "node_order": f"top {levels} levels ({M} nodes) breadth-first at the front", "ms": ms, "mrays_s": R / ms / 1e3, "identical_results": same}), flush=True)
        cs2.close()


if __name__ == "__main__":
    main()
