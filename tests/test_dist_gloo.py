"""CPU, world_size 2 over gloo: the N>1 host logic of engine_b200/dist.py — ray sharding bounds, BVH
blob replication and the ordered gather of results.  The tracer is the oracle port here (test
infrastructure; on the GPU box the same functions are handed CollisionSystem methods)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_rays, q):
    import sys

    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from engine_b200 import dist as b2dist
        from engine_b200 import synth
        from oracle.oracle import NODE_DTYPE, RAYCAST_DTYPE, Port

        port_lib = Port()
        cpu = torch.device("cpu")
        # rank 0 "builds" a tree; the blob that travels is its byte image
        state = {}
        if rank == 0:
            s = synth.uniform_boxes(3000, seed=7, extent=50)
            world_m = port_lib.transform_update(s.parent, s.pos, s.rot, s.scl)
            boxes = port_lib.refit(s.collider_entity, s.collider_aabb, world_m)
            state["tree"] = port_lib.build_binned_sah(s.collider_entity, boxes)

        def export_fn():
            return torch.from_numpy(state["tree"].view(np.uint8).reshape(-1).copy())

        def import_fn(t):
            state["tree"] = t.numpy().view(NODE_DTYPE).reshape(-1).copy()

        nbytes = b2dist.replicate_blob(export_fn if rank == 0 else None, import_fn, cpu, src=0)
        assert nbytes == state["tree"].nbytes
        rays = synth.random_rays(n_rays, seed=8, extent=60)
        lo, hi = b2dist.shard_bounds(n_rays, world, rank)
        traced = []

        def trace_fn(r):
            traced.append(len(r))
            return port_lib.raycast_tree(state["tree"], r)

        out = b2dist.raycast_sharded(trace_fn, rays, RAYCAST_DTYPE, cpu, dst=0)
        assert traced == [hi - lo]
        # device-resident flavour of the gather (here: CPU tensors over gloo), equal-sized shards
        even = n_rays - n_rays % world
        lo2, hi2 = b2dist.shard_bounds(even, world, rank)
        mine = torch.from_numpy(port_lib.raycast_tree(state["tree"], rays[lo2:hi2]).view(np.uint8).reshape(-1).copy())
        gathered = b2dist.gather_hits_device(mine, dst=0)
        if rank == 0:
            want = port_lib.raycast_tree(state["tree"], rays[:even]).view(np.uint8).reshape(-1)
            assert (gathered.numpy() == want).all()
        else:
            assert gathered is None
        if rank == 0:
            whole = port_lib.raycast_tree(state["tree"], rays)
            q.put(bool((out.view(np.uint8) == whole.view(np.uint8)).all()) and int(whole["hit"].sum()) > 0)
        else:
            assert out is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_rays", [1000, 1001])
def test_sharded_raycast_equals_unsharded_world2(n_rays):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_rays, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_shard_bounds_cover_batch():
    from engine_b200.dist import shard_bounds

    for n in (0, 1, 7, 16, 1_000_003):
        for world in (1, 2, 3, 8):
            edges = [shard_bounds(n, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1
