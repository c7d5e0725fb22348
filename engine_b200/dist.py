"""Multi-GPU host logic: one process per GPU, torch.distributed for the plumbing (NCCL on the GPU
box, gloo in the CPU tests).  The path shards by RAYS; the BVH is replicated:

  rank `src` builds (b2col_update) -> one broadcast of the BVH blob (NVLink) -> every rank imports it
  -> each rank traces its contiguous slice of the batch -> slices are gathered back in order.

There is no collective on the per-ray data path; the only exchanges are the blob broadcast (once per
rebuild) and the optional gather of results.  The functions take the tracer / blob accessors as
callables so that the sharding and collective logic can be exercised without a GPU
(tests/test_dist_gloo.py drives them with world_size 2 over gloo).
"""
from __future__ import annotations

from typing import Callable

import numpy as np
import torch
import torch.distributed as dist


def shard_bounds(n: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous, order-preserving slice [lo, hi) of an n-ray batch for `rank`."""
    return n * rank // world, n * (rank + 1) // world


def replicate_blob(export_fn: Callable[[], torch.Tensor] | None, import_fn: Callable[[torch.Tensor], None], device, src: int = 0, group=None) -> int:
    """Broadcast the BVH blob from `src` to every rank. export_fn (only called on src) returns a uint8
    tensor on `device`; import_fn (called on the other ranks) consumes the received tensor.
    Returns the blob size in bytes."""
    rank = dist.get_rank(group)
    size = torch.zeros(1, dtype=torch.int64, device=device)
    blob = None
    if rank == src:
        blob = export_fn()
        size[0] = blob.numel()
    dist.broadcast(size, src, group=group)
    nbytes = int(size.item())
    if rank != src:
        blob = torch.empty(nbytes, dtype=torch.uint8, device=device)
    dist.broadcast(blob, src, group=group)
    if rank != src:
        import_fn(blob)
    return nbytes


def replicate_bvh(cs, src: int = 0, group=None) -> int:
    """CollisionSystem flavour of replicate_blob (CUDA tensors, NCCL)."""
    device = torch.device("cuda", cs.device)

    def export_fn():
        _, nbytes = cs.bvh_blob()
        t = torch.empty(nbytes, dtype=torch.uint8, device=device)
        torch.cuda.synchronize(device)
        cs.bvh_export_blob(t, nbytes)
        cs.sync()
        return t

    def import_fn(t):
        # the broadcast ran on torch's stream; the import copies on the context's own: order them here
        torch.cuda.current_stream(device).synchronize()
        cs.bvh_import_blob(t, t.numel())
        cs.sync()

    return replicate_blob(export_fn, import_fn, device, src, group)


def raycast_sharded(trace_fn: Callable[[np.ndarray], np.ndarray], rays: np.ndarray, result_dtype, device, dst: int = 0, group=None):
    """Every rank holds the same `rays` (n,6); rank r traces slice shard_bounds(n, world, r) with
    trace_fn and the slices are gathered, in order, on `dst` (returns None elsewhere).  The output is
    byte-identical to tracing the whole batch on one GPU because rays are independent."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    n = rays.shape[0]
    lo, hi = shard_bounds(n, world, rank)
    mine = np.ascontiguousarray(trace_fn(rays[lo:hi]))
    item = np.dtype(result_dtype).itemsize
    t = torch.from_numpy(mine.view(np.uint8).reshape(-1)).to(device)
    # slices may differ by one ray: pad to the largest so that gather sees equal shapes
    longest = max(shard_bounds(n, world, r)[1] - shard_bounds(n, world, r)[0] for r in range(world)) * item
    padded = torch.zeros(longest, dtype=torch.uint8, device=device)
    padded[: t.numel()] = t
    parts = [torch.empty_like(padded) for _ in range(world)] if rank == dst else None
    dist.gather(padded, parts, dst=dst, group=group)
    if rank != dst:
        return None
    out = np.empty(n, dtype=result_dtype)
    ob = out.view(np.uint8).reshape(-1)
    for r in range(world):
        a, b = shard_bounds(n, world, r)
        ob[a * item : b * item] = parts[r][: (b - a) * item].cpu().numpy()
    return out


def gather_hits_device(d_out: torch.Tensor, dst: int = 0, group=None, into: torch.Tensor | None = None):
    """Device-resident flavour of the result gather: every rank's Raycast bytes (equal sizes) land, in rank order, in
    one tensor on `dst` (NCCL gather over NVLink; `into` lets the caller reuse the world x n destination).  Returns the
    gathered tensor on dst, None elsewhere.  Shards are contiguous slices of the batch, so the concatenation IS the
    unsharded result."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    parts = None
    if rank == dst:
        if into is None:
            into = torch.empty(world * d_out.numel(), dtype=d_out.dtype, device=d_out.device)
        parts = list(into.view(world, -1).unbind(0))
    dist.gather(d_out, parts, dst=dst, group=group)
    return into if rank == dst else None


def bind_to_gpu_numa_node(device_index: int) -> str | None:
    """Pin this process (and the pinned host buffers it allocates afterwards) to the CPUs of the NUMA node its GPU hangs
    off — what `numactl --cpunodebind` would do per rank.  With 8 ranks on a two-socket host, staging buffers on the
    far socket halve the PCIe copy rate.  Best effort: returns the cpulist it bound to, or None."""
    try:
        import os

        import pynvml

        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        index = int(visible.split(",")[device_index]) if visible and all(v.strip().isdigit() for v in visible.split(",")) else device_index
        bus_id = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus_id = bus_id.decode() if isinstance(bus_id, bytes) else bus_id
        domain, rest = bus_id.lower().split(":", 1)
        pci = f"{domain[-4:]}:{rest}"
        node = int(open(f"/sys/bus/pci/devices/{pci}/numa_node").read())
        if node < 0:
            return None
        cpulist = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
        cpus = set()
        for part in cpulist.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return cpulist
    except Exception:
        return None
