"""Host-side mirror of the reference's collision interface over the C ABI (include/b2col.h).

``CollisionSystem`` keeps the reference class's vocabulary (reference include/system_collisions.h:28-74):
``on_update`` = TransformSystem::onUpdate + CollisionSystem::onUpdate (src/system_transform.cpp:12-38,
src/system_collisions.cpp:112-158), ``get_raycast`` = GetRaycast (:160-199), ``bvh`` = the public
``bvh_`` member (:58).  ``get_raycasts`` is the additive batch entry.  All compute happens in
libb2col.so on the GPU; this module only marshals numpy / torch buffers.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .capi import AABB_DTYPE, MODE_LBVH, MODE_REFTOPO, NODE_DTYPE, RAY_DTYPE, RAYCAST_DTYPE, B2ColError, ptr


def _rays(rays) -> np.ndarray:
    rays = np.ascontiguousarray(rays)
    if rays.dtype != RAY_DTYPE:
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
    return rays


class CollisionSystem:
    def __init__(self, device: int = 0, max_distance: float = 1000.0, morton_bits_per_axis: int = 0, use_cuda_graph: bool = False):
        self.lib = capi.load()
        cfg = capi.Config(device=device, max_distance=max_distance, morton_bits_per_axis=morton_bits_per_axis, use_cuda_graph=int(use_cuda_graph))
        h = C.c_void_p()
        st = self.lib.b2col_create(C.byref(h), C.byref(cfg))
        if st != capi.OK:
            raise B2ColError(st, (self.lib.b2col_last_error(None) or b"").decode())
        self.h = h
        self.device = device
        self.n_entities = 0
        self.n_colliders = 0

    # -- plumbing ------------------------------------------------------------------------------
    def _ck(self, st: int) -> None:
        if st != capi.OK:
            raise B2ColError(st, (self.lib.b2col_last_error(self.h) or b"").decode())

    def close(self) -> None:
        if getattr(self, "h", None):
            self.lib.b2col_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self) -> None:
        self._ck(self.lib.b2col_sync(self.h))

    @property
    def stream(self) -> int:
        return int(self.lib.b2col_stream(self.h) or 0)

    # -- scene input ---------------------------------------------------------------------------
    def set_scene(self, scene) -> None:
        """scene: engine_b200.synth.Scene (arrays indexed by entity ID, row 0 unused)."""
        self.set_hierarchy(scene.parent)
        self.set_trs(scene.pos, scene.rot, scene.scl)
        self.set_colliders(scene.collider_entity, scene.collider_aabb)

    def set_hierarchy(self, parent) -> None:
        parent = np.ascontiguousarray(parent, dtype=np.uint32)
        self.n_entities = parent.shape[0] - 1
        self._ck(self.lib.b2col_set_hierarchy(self.h, self.n_entities, ptr(parent)))

    def set_trs(self, pos, rot_xyzw, scl) -> None:
        n = self.n_entities
        pos = np.ascontiguousarray(pos, dtype=np.float32).reshape(n + 1, 3)
        rot = np.ascontiguousarray(rot_xyzw, dtype=np.float32).reshape(n + 1, 4)
        scl = np.ascontiguousarray(scl, dtype=np.float32).reshape(n + 1, 3)
        self._ck(self.lib.b2col_set_trs(self.h, n, ptr(pos), ptr(rot), ptr(scl)))

    def update_trs_rows(self, entity, pos, rot_xyzw, scl) -> None:
        """sparse update: rows entity[i] of the device TRS arrays take pos / rot / scl [i] (b2col_update_trs_rows)"""
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        k = entity.shape[0]
        pos = np.ascontiguousarray(pos, dtype=np.float32).reshape(k, 3)
        rot = np.ascontiguousarray(rot_xyzw, dtype=np.float32).reshape(k, 4)
        scl = np.ascontiguousarray(scl, dtype=np.float32).reshape(k, 3)
        self._ck(self.lib.b2col_update_trs_rows(self.h, k, ptr(entity), ptr(pos), ptr(rot), ptr(scl)))

    def world_rows(self, entity) -> np.ndarray:
        """world matrices of the given entities, gathered on the device (b2col_fetch_world_rows)"""
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        out = C.c_void_p()
        self._ck(self.lib.b2col_fetch_world_rows(self.h, entity.shape[0], ptr(entity), C.byref(out)))
        if entity.shape[0] == 0:
            return np.zeros((0, 16), np.float32)
        buf = (C.c_float * (16 * entity.shape[0])).from_address(out.value)
        return np.frombuffer(buf, dtype=np.float32).reshape(-1, 16).copy()

    def set_trs_device(self, d_pos, d_rot_xyzw, d_scl) -> None:
        """torch CUDA tensors (float32, contiguous, (E+1,3)/(E+1,4)/(E+1,3)); enqueued, not waited for."""
        self._ck(self.lib.b2col_set_trs_device(self.h, self.n_entities, ptr(d_pos), ptr(d_rot_xyzw), ptr(d_scl)))

    def bind_trs_device(self, d_pos, d_rot_xyzw, d_scl) -> None:
        """zero-copy: the context reads these torch CUDA tensors in place at every update until the next set_trs* (keep them alive;
        order your writes on self.stream or with events)"""
        self._bound_trs = (d_pos, d_rot_xyzw, d_scl)  # keep the tensors alive as long as the binding
        self._ck(self.lib.b2col_bind_trs_device(self.h, self.n_entities, ptr(d_pos), ptr(d_rot_xyzw), ptr(d_scl)))

    def set_world_matrices(self, world16) -> None:
        world16 = np.ascontiguousarray(world16, dtype=np.float32).reshape(-1, 16)
        self.n_entities = world16.shape[0] - 1
        self._ck(self.lib.b2col_set_world_matrices(self.h, self.n_entities, ptr(world16)))

    def set_colliders(self, entity, local_aabb) -> None:
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        local = np.ascontiguousarray(local_aabb)
        if local.dtype != AABB_DTYPE:
            local = np.ascontiguousarray(local, dtype=np.float32).reshape(-1, 6)
        self.n_colliders = entity.shape[0]
        self._ck(self.lib.b2col_set_colliders(self.h, self.n_colliders, ptr(entity), ptr(local)))

    # -- per-frame -----------------------------------------------------------------------------
    def on_update(self, ts: float = 0.0, wait: bool = True) -> None:
        self._ck(self.lib.b2col_update(self.h))
        if wait:
            self.sync()

    def on_update_refit(self, ts: float = 0.0, wait: bool = True) -> None:
        """Refit-only frame (the reference's TODO at src/system_collisions.cpp:116-120): same tree shape, new boxes."""
        self._ck(self.lib.b2col_refit(self.h))
        if wait:
            self.sync()

    # -- queries -------------------------------------------------------------------------------
    def propagate(self, wait: bool = True) -> None:
        """TransformSystem::onUpdate alone (reference src/system_transform.cpp:12-38): world matrices on the device."""
        self._ck(self.lib.b2col_propagate(self.h))
        if wait:
            self.sync()

    def get_raycasts(self, rays, mode: int = MODE_LBVH, out: np.ndarray | None = None) -> np.ndarray:
        rays = _rays(rays)
        n = rays.shape[0]
        if out is None:
            out = np.empty(n, RAYCAST_DTYPE)
        self._ck(self.lib.b2col_raycast_batch(self.h, ptr(rays), n, ptr(out), mode))
        return out

    def get_hits(self, rays, out: np.ndarray | None = None) -> np.ndarray:
        """compact results (b2col_hit: entity, distance, hit | side << 8): 12 instead of 36 bytes per ray leave the device"""
        rays = _rays(rays)
        n = rays.shape[0]
        if out is None:
            out = np.empty(n, capi.HIT_DTYPE)
        self._ck(self.lib.b2col_raycast_hits_batch(self.h, ptr(rays), n, ptr(out)))
        return out

    def expand_hits(self, rays, hits, threads: int = 1, out: np.ndarray | None = None) -> np.ndarray:
        """b2col_expand_hits: the Raycasts of get_raycasts from the compact hits, computed on the host"""
        rays = _rays(rays)
        n = rays.shape[0]
        if out is None:
            out = np.empty(n, RAYCAST_DTYPE)
        st = self.lib.b2col_expand_hits(ptr(rays), ptr(hits), n, ptr(out), threads)
        assert st == capi.OK
        return out

    def raycast_hits_device(self, d_rays, n: int, d_out) -> None:
        self._ck(self.lib.b2col_raycast_hits_device(self.h, ptr(d_rays), n, ptr(d_out)))

    def raycast_hits_host_pinned(self, h_rays, n: int, h_out) -> None:
        self._ck(self.lib.b2col_raycast_hits_batch(self.h, ptr(h_rays), n, ptr(h_out)))

    def get_raycast(self, origin, direction, mode: int = MODE_LBVH):
        ray = np.zeros((1, 6), np.float32)
        ray[0, :3] = origin
        ray[0, 3:] = direction
        return self.get_raycasts(ray, mode)[0]

    def raycast_device(self, d_rays, n: int, d_out, mode: int = MODE_LBVH) -> None:
        """Device-resident rays/results (torch CUDA tensors or raw addresses); enqueued on the ctx stream."""
        self._ck(self.lib.b2col_raycast_device(self.h, ptr(d_rays), n, ptr(d_out), mode))

    def raycast_host_pinned(self, h_rays, n: int, h_out, mode: int = MODE_LBVH) -> None:
        """Host buffers given as torch pinned tensors / raw addresses (the e2e path of bench.py)."""
        self._ck(self.lib.b2col_raycast_batch(self.h, ptr(h_rays), n, ptr(h_out), mode))

    # -- reference-topology mode / debug overlay -------------------------------------------------
    def upload_ref_bvh(self, nodes) -> None:
        nodes = np.ascontiguousarray(nodes, dtype=NODE_DTYPE)
        self._ck(self.lib.b2col_upload_ref_bvh(self.h, ptr(nodes), nodes.shape[0]))

    @property
    def bvh(self) -> np.ndarray:
        """The GPU-built tree in the reference's BiTreeNode format (root last)."""
        cap = max(2 * self.n_colliders, 1)
        out = np.zeros(cap, NODE_DTYPE)
        n = C.c_size_t(0)
        self._ck(self.lib.b2col_export_bitreenode(self.h, ptr(out), cap, C.byref(n)))
        return out[: n.value].copy()

    # -- read-back -----------------------------------------------------------------------------
    def world_matrices(self, first: int = 1, count: int | None = None) -> np.ndarray:
        count = self.n_entities - first + 1 if count is None else count
        out = np.empty((count, 16), np.float32)
        self._ck(self.lib.b2col_get_world_matrices(self.h, first, count, ptr(out)))
        return out

    def world_aabbs(self) -> np.ndarray:
        out = np.empty((self.n_colliders, 6), np.float32)
        self._ck(self.lib.b2col_get_world_aabbs(self.h, ptr(out)))
        return out

    def read_raycasts(self, d_addr: int, n: int) -> np.ndarray:
        """n Raycasts from device memory of this context (e.g. the gathered array of comm_raycast_gather), after everything enqueued"""
        out = np.empty(n, RAYCAST_DTYPE)
        self._ck(self.lib.b2col_read_device(self.h, C.c_void_p(d_addr), n * RAYCAST_DTYPE.itemsize, ptr(out)))
        return out

    def leaf_order(self) -> np.ndarray:
        out = np.empty(self.n_colliders, np.uint32)
        self._ck(self.lib.b2col_get_leaf_order(self.h, ptr(out)))
        return out

    def stats(self) -> dict:
        s = capi.Stats()
        self._ck(self.lib.b2col_get_stats(self.h, C.byref(s)))
        d = {k: getattr(s, k) for k, _ in capi.Stats._fields_ if k not in ("scene_min", "scene_max")}
        d["scene_min"], d["scene_max"] = list(s.scene_min), list(s.scene_max)
        return d

    def trace_counters(self) -> dict:
        """totals over all MODE_LBVH_COUNT calls: BVH nodes fetched (one 32-byte sector each) and rays traced"""
        v, r = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.b2col_get_trace_counters(self.h, C.byref(v), C.byref(r)))
        return {"node_visits": int(v.value), "rays": int(r.value)}

    # -- multi-GPU (include/b2col.h "multi-GPU"; NCCL is called from C++ inside libb2col.so) ------
    @staticmethod
    def comm_unique_id() -> bytes:
        """ncclGetUniqueId (rank 0 calls it; distribute the 128 bytes by any means)"""
        lib = capi.load()
        buf = C.create_string_buffer(capi.COMM_ID_BYTES)
        st = lib.b2col_comm_get_unique_id(buf)
        if st != capi.OK:
            raise B2ColError(st, (lib.b2col_last_error(None) or b"").decode())
        return buf.raw

    def comm_init_rank(self, n_ranks: int, rank: int, unique_id: bytes) -> None:
        assert len(unique_id) == capi.COMM_ID_BYTES
        self._ck(self.lib.b2col_comm_init_rank(self.h, n_ranks, rank, C.c_char_p(unique_id)))

    def comm_finalize(self) -> None:
        self._ck(self.lib.b2col_comm_finalize(self.h))

    def comm_replicate(self, src_rank: int, n_colliders: int) -> None:
        """collective: one ncclBroadcast of rank src's BVH blob into every rank's blob allocation (enqueued)"""
        self._ck(self.lib.b2col_comm_replicate(self.h, src_rank, n_colliders))

    def comm_replicate_records(self, src_rank: int, n_colliders: int) -> None:
        """collective: broadcast the 32-byte world records instead and build the (identical) tree on every rank"""
        self._ck(self.lib.b2col_comm_replicate_records(self.h, src_rank, n_colliders))

    def comm_raycast_gather(self, d_rays_shard, n_total: int, dst_rank: int = 0) -> int:
        """collective: trace this rank's slice of an n_total-ray batch; the Raycasts of the whole batch land in batch order on
        dst_rank.  Returns the device address of the gathered array on dst_rank, 0 elsewhere.  Enqueued."""
        out = C.c_void_p()
        self._ck(self.lib.b2col_comm_raycast_gather(self.h, ptr(d_rays_shard), n_total, dst_rank, C.byref(out)))
        return int(out.value or 0)

    def comm_info(self) -> dict:
        ci = capi.CommInfo()
        self._ck(self.lib.b2col_comm_info(self.h, C.byref(ci)))
        return {k: getattr(ci, k) for k, _ in capi.CommInfo._fields_}

    def bvh_blob(self) -> tuple[int, int]:
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self.lib.b2col_bvh_blob(self.h, C.byref(p), C.byref(n)))
        return int(p.value or 0), int(n.value)

    def bvh_export_blob(self, d_dst, nbytes: int) -> None:
        self._ck(self.lib.b2col_bvh_export_blob(self.h, ptr(d_dst), nbytes))

    def bvh_import_blob(self, d_src, nbytes: int) -> None:
        self._ck(self.lib.b2col_bvh_import_blob(self.h, ptr(d_src), nbytes))

    # -- timing --------------------------------------------------------------------------------
    def event_record(self, slot: int) -> None:
        self._ck(self.lib.b2col_event_record(self.h, slot))

    def event_elapsed_ms(self, a: int, b: int) -> float:
        ms = C.c_float()
        self._ck(self.lib.b2col_event_elapsed_ms(self.h, a, b, C.byref(ms)))
        return float(ms.value)


def shard_bounds(n: int, n_ranks: int, rank: int) -> tuple[int, int]:
    """b2col_shard_bounds: rank's contiguous slice [lo, hi) of an n-element batch"""
    lo, hi = C.c_size_t(), C.c_size_t()
    capi.load().b2col_shard_bounds(n, n_ranks, rank, C.byref(lo), C.byref(hi))
    return int(lo.value), int(hi.value)


class MultiGpu:
    """Several GPUs driven by ONE process (how a game engine would use the backend; engine_b200/shim does the same in
    C++): context 0 owns the scene and builds, the others hold replicas.  b2col_comm_init_all / _replicate_all /
    _raycast_gather_all / b2col_mgpu_raycast_batch."""

    def __init__(self, devices, **kw):
        self.systems = [CollisionSystem(device=d, **kw) for d in devices]
        self.lib = self.systems[0].lib
        self._arr = (C.c_void_p * len(self.systems))(*[cs.h for cs in self.systems])
        self.primary._ck(self.lib.b2col_comm_init_all(self._arr, len(self.systems)))

    @property
    def primary(self) -> CollisionSystem:
        return self.systems[0]

    def replicate(self, src: int = 0) -> None:
        self.primary._ck(self.lib.b2col_comm_replicate_all(self._arr, len(self.systems), src))

    def raycast_gather(self, d_rays_shards, n_total: int, dst: int = 0) -> int:
        """d_rays_shards[i]: device tensor / address of rank i's slice (on device i). Returns the device address (on dst) of the
        gathered Raycasts; enqueued — sync the dst system before reading."""
        shards = (C.c_void_p * len(self.systems))(*[ptr(t) for t in d_rays_shards])
        out = C.c_void_p()
        self.primary._ck(self.lib.b2col_comm_raycast_gather_all(self._arr, len(self.systems), shards, n_total, dst, C.byref(out)))
        return int(out.value or 0)

    def get_raycasts(self, rays, out: np.ndarray | None = None) -> np.ndarray:
        rays = _rays(rays)
        n = rays.shape[0]
        if out is None:
            out = np.empty(n, RAYCAST_DTYPE)
        self.primary._ck(self.lib.b2col_mgpu_raycast_batch(self._arr, len(self.systems), ptr(rays), n, ptr(out), MODE_LBVH))
        return out

    def raycast_host_pinned(self, h_rays, n: int, h_out) -> None:
        self.primary._ck(self.lib.b2col_mgpu_raycast_batch(self._arr, len(self.systems), ptr(h_rays), n, ptr(h_out), MODE_LBVH))

    def close(self) -> None:
        for cs in self.systems:
            cs.close()
