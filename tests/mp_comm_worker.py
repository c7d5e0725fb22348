"""TEST INFRASTRUCTURE — one rank of the multi-process (one process per GPU) data plane of libb2col.so, launched by
tests/test_gpu_parity.py::test_multi_gpu_one_process_per_gpu with torch.distributed.run.  torch.distributed (gloo) only carries
the 128-byte NCCL unique id and a final verdict; the broadcast of the BVH and the gather of the hits are b2col_comm_* calls, i.e.
NCCL / peer stores issued from C++ inside the library."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from engine_b200 import synth  # noqa: E402
from engine_b200.collision import CollisionSystem, shard_bounds  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    n_colliders, n_rays = int(sys.argv[1]), int(sys.argv[2])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    ids = [CollisionSystem.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    cs = CollisionSystem(device=local, use_cuda_graph=True)
    cs.comm_init_rank(world, rank, ids[0])
    scene = synth.uniform_boxes(n_colliders, seed=21, extent=400.0)
    rays = synth.random_rays(n_rays, seed=22, extent=420.0)  # every rank generates the whole batch, traces its slice
    lo, hi = shard_bounds(n_rays, world, rank)
    d_shard = torch.from_numpy(rays[lo:hi]).cuda()
    ok = True
    for how in ("blob", "records", "blob"):
        if rank == 0:
            if how != "blob":
                synth_move = scene.pos.copy()
                synth_move[1:] += np.float32(0.25)
                cs.set_hierarchy(scene.parent)
                cs.set_trs(synth_move, scene.rot, scene.scl)
                cs.set_colliders(scene.collider_entity, scene.collider_aabb)
            else:
                cs.set_scene(scene)
            cs.on_update()
        if how == "blob":
            cs.comm_replicate(0, n_colliders)
        else:
            cs.comm_replicate_records(0, n_colliders)
        for dst in (0, world - 1):
            addr = cs.comm_raycast_gather(d_shard, n_rays, dst)
            cs.sync()
            dist.barrier()  # every rank's kernel has completed (the library's own completion step already ordered dst's stream)
            if rank == dst:
                got = cs.read_raycasts(addr, n_rays)
            if rank == 0:
                want = cs.get_raycasts(rays)  # the 1-GPU trace of the same batch, on the rank that built the tree
            verdict = [None]
            if dst != 0:
                box = [got if rank == dst else None]
                dist.broadcast_object_list(box, src=dst)
                got = box[0]
            if rank == 0:
                same = bool((got.view(np.uint8) == want.view(np.uint8)).all())
                ok = ok and same and want["hit"].sum() > 100
                print(f"[mp_comm_worker] replicate={how} dst={dst} gather_mode={cs.comm_info()['gather_mode']} identical={same} hits={int(want['hit'].sum())}", flush=True)
    flag = [ok]
    dist.broadcast_object_list(flag, src=0)
    info = cs.comm_info()
    cs.comm_finalize()
    cs.close()
    dist.destroy_process_group()
    if rank == 0:
        print("MP_COMM_OK" if flag[0] else "MP_COMM_MISMATCH", info, flush=True)
    sys.exit(0 if flag[0] else 1)


if __name__ == "__main__":
    main()
