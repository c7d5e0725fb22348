#!/usr/bin/env python
"""Dev experiment (not product): how much would a higher-quality tree buy k_trace_lbvh?

The production builder is an LBVH (Morton order).  This script builds the same scene's tree with the oracle's
binned-SAH builder on the HOST (test infrastructure, oracle/port.c), re-expresses it in the traversal blob's
format (QNode / LeafRec, leaves numbered in depth-first order so the tie rule is unchanged), imports it with
b2col_bvh_import_blob and times the UNCHANGED trace kernel on both trees.  It answers "is a GPU SAH/PLOC builder
worth writing for static scenes?" with a measurement instead of a guess.

  python tools/exp_tree_quality.py --colliders 1000000 --rays 16777216 --workload config3
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

LEAF_BIT = 0x80000000
GRID_CELLS = 32768.0


def quant(x, gmin, inv_cell, hi):
    f = (x.astype(np.float32) - gmin) * inv_cell  # float32, no FMA: same ops as quant_lo / quant_hi
    f = (np.ceil(f) + np.float32(2.0)) if hi else (np.floor(f) - np.float32(2.0))
    f = np.clip(f, 0.0, GRID_CELLS - 1.0)
    return (np.uint32(0x8000) | f.astype(np.uint32)).astype(np.uint32)


def convert(nodes, header_words, entity_to_prim):
    """reference-shaped tree (root = last node) -> (root index, qnodes [n-1, 8] u32, leaf records [n, 8] u32)"""
    from oracle.oracle import TYPE_BV, TYPE_ENTITY, TYPE_EMPTY

    grid = header_words[16:25].view(np.float32)
    gmin, inv_cell = grid[0:3].copy(), grid[6:9].copy()
    t1, t2 = nodes["type1"], nodes["type2"]
    i1, i2 = nodes["index1"], nodes["index2"]
    b1 = np.concatenate([nodes["box1"]["min"], nodes["box1"]["max"]], axis=1).astype(np.float32)
    b2 = np.concatenate([nodes["box2"]["min"], nodes["box2"]["max"]], axis=1).astype(np.float32)
    wrapper = (t1 == TYPE_ENTITY) & (t2 == TYPE_EMPTY)  # {Entity, Empty}: a single leaf wrapped in a node
    M = nodes.shape[0]
    new_id = np.full(M, -1, np.int64)
    # depth-first walk (slot 1 before slot 2): internal ids in visit order, leaf positions in DFS order
    leaf_ent, leaf_box, refs = [], [], np.zeros((M, 2), np.uint32)
    n_int = 0
    stack = [M - 1]
    # iterative DFS that emits leaves in order: push slot 2 then slot 1
    order = []
    work = [("node", M - 1)]
    while work:
        kind, v = work.pop()
        if kind == "leaf":
            order.append(v)
            continue
        if wrapper[v]:
            order.append((v, 0))
            continue
        new_id[v] = n_int
        n_int += 1
        # slot 2 first on the stack so slot 1 is processed first
        if t2[v] == TYPE_BV:
            work.append(("node", int(i2[v])))
        elif t2[v] == TYPE_ENTITY:
            work.append(("leaf", (v, 1)))
        if t1[v] == TYPE_BV:
            work.append(("node", int(i1[v])))
        elif t1[v] == TYPE_ENTITY:
            work.append(("leaf", (v, 0)))
    n = len(order)
    assert n_int == n - 1, (n_int, n)
    pos_of = {}
    lf = np.zeros((n, 8), np.uint32)
    for p, (v, s) in enumerate(order):
        pos_of[(v, s)] = p
        box = b1[v] if s == 0 else b2[v]
        ent = int(i1[v]) if s == 0 else int(i2[v])
        lf[p, 0:6] = box.view(np.uint32)
        lf[p, 6] = ent
        lf[p, 7] = entity_to_prim[ent]
    q = np.zeros((n - 1, 8), np.uint32)
    iv = np.nonzero(new_id >= 0)[0]
    for slot, (tt, ii, bb) in enumerate(((t1, i1, b1), (t2, i2, b2))):
        box = bb[iv].copy()
        ref = np.zeros(iv.shape[0], np.uint32)
        for k, v in enumerate(iv):
            if tt[v] == TYPE_BV:
                c = int(ii[v])
                if wrapper[c]:
                    ref[k] = LEAF_BIT | pos_of[(c, 0)]
                    box[k] = b1[c]
                else:
                    ref[k] = new_id[c]
            else:
                assert tt[v] == TYPE_ENTITY
                ref[k] = LEAF_BIT | pos_of[(int(v), slot)]
        for ax in range(3):
            q[new_id[iv], 3 * slot + ax] = quant(box[:, ax], gmin[ax], inv_cell[ax], False) | (quant(box[:, 3 + ax], gmin[ax], inv_cell[ax], True) << np.uint32(16))
        q[new_id[iv], 6 + slot] = ref
    return int(new_id[M - 1]), q, lf


def sah_cost(q, root):
    """sum of the surface areas of all child boxes (grid cells^2) ~ expected node visits of long random rays"""
    lo = (q[:, 0:6] & 0x7FFF).astype(np.float64)
    hi = ((q[:, 0:6] >> 16) & 0x7FFF).astype(np.float64)
    e = hi - lo
    a = 0.0
    for s in (0, 3):
        a += (e[:, s] * e[:, s + 1] + e[:, s + 1] * e[:, s + 2] + e[:, s] * e[:, s + 2]).sum()
    return a


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--colliders", type=int, default=1_000_000)
    ap.add_argument("--rays", type=int, default=1 << 24)
    ap.add_argument("--workload", default="config3", choices=["config3", "config5", "config2"])
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    import torch

    from engine_b200 import synth
    from engine_b200.collision import CollisionSystem
    from oracle.oracle import Port

    if args.workload == "config3":
        s = synth.uniform_boxes(args.colliders)
        rays = synth.random_rays(args.rays)
    elif args.workload == "config5":
        s = synth.clustered_boxes(args.colliders)
        rays = synth.random_rays(args.rays, seed=4, extent=2000.0)
    else:
        s = synth.static_city(args.colliders)
        w = int(np.sqrt(args.rays))
        rays = synth.camera_rays(w, w)
    R = rays.shape[0]
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
        for _ in range(args.steps):
            c.raycast_device(d_rays, R, d_out)
        c.event_record(3)
        c.sync()
        return c.event_elapsed_ms(2, 3) / args.steps

    ms_lbvh = time_trace(cs)
    res_lbvh = d_out.cpu().numpy().reshape(R, 36).copy()
    _, nbytes = cs.bvh_blob()
    blob = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    cs.bvh_export_blob(blob, nbytes)
    cs.sync()
    words = blob.cpu().numpy().view(np.uint32).copy()
    n = int(words[1])
    q_lbvh = words[64 : 64 + 8 * (n - 1)].reshape(n - 1, 8)

    port = Port()
    boxes = cs.world_aabbs()
    t0 = time.perf_counter()
    tree = port.build_binned_sah(s.collider_entity, boxes)
    t_build = time.perf_counter() - t0
    e2p = np.zeros(int(s.collider_entity.max()) + 1, np.uint32)
    e2p[s.collider_entity] = np.arange(s.n_colliders, dtype=np.uint32)
    t0 = time.perf_counter()
    root, q, lf = convert(tree, words[:64], e2p)
    t_conv = time.perf_counter() - t0
    out = words.copy()
    out[3] = root
    out[25], out[26] = root, 0  # trace_root, top_count (no compact top copy)
    out[64 : 64 + 8 * (n - 1)] = q.reshape(-1)
    out[64 + 8 * (n - 1) : 64 + 8 * (n - 1) + 8 * n] = lf.reshape(-1)
    blob2 = torch.from_numpy(out.view(np.uint8)).cuda()
    torch.cuda.synchronize()
    cs2 = CollisionSystem(device=0)
    cs2.bvh_import_blob(blob2, nbytes)
    ms_sah = time_trace(cs2)
    res_sah = d_out.cpu().numpy().reshape(R, 36)
    hit_same = (res_lbvh[:, 32] == res_sah[:, 32]).mean()
    ent_same = (res_lbvh[:, 24:28] == res_sah[:, 24:28]).all(1).mean()
    dist_same = (res_lbvh[:, 28:32] == res_sah[:, 28:32]).all(1).mean()
    print(json.dumps({
        "workload": args.workload, "colliders": n, "rays": R,
        "lbvh": {"ms": ms_lbvh, "mrays_s": R / ms_lbvh / 1e3, "sa_cost": sah_cost(q_lbvh, int(words[3]))},
        "binned_sah": {"ms": ms_sah, "mrays_s": R / ms_sah / 1e3, "sa_cost": sah_cost(q, root), "host_build_s": t_build, "convert_s": t_conv},
        "speedup": ms_lbvh / ms_sah, "hit_flag_equal": hit_same, "entity_equal": ent_same, "distance_equal": dist_same,
    }))


if __name__ == "__main__":
    main()
