"""ctypes binding of include/b2col.h (libb2col.so).  No CPU path: loading fails loudly when the
library has not been built, and ``b2col_create`` fails with B2COL_ERR_NO_DEVICE without a GPU."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2COL_LIB") or os.path.join(HERE, "libb2col.so")  # B2COL_LIB: dev override (kernel variants)

OK, ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_STATE, ERR_UNSUPPORTED, ERR_OVERFLOW, ERR_COMM = range(8)
COMM_ID_BYTES = 128
GATHER_NONE, GATHER_PEER_WINDOW, GATHER_PEER_WINDOW_IPC, GATHER_NCCL_SENDRECV = range(4)
MODE_LBVH, MODE_REFTOPO, MODE_LBVH_COUNT = 0, 1, 2

# PODs — byte-identical to the reference's (include/system_collisions.h:13-24,40-48)
RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("direction", "<f4", 3)])
RAYCAST_DTYPE = np.dtype(
    [("location", "<f4", 3), ("normal", "<f4", 3), ("hit_entity", "<u4"), ("distance", "<f4"), ("hit", "u1"), ("pad", "u1", 3)]
)
HIT_DTYPE = np.dtype([("hit_entity", "<u4"), ("distance", "<f4"), ("flags", "<u4")])  # b2col_hit
AABB_DTYPE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3)])
NODE_DTYPE = np.dtype(
    [("box1", AABB_DTYPE), ("box2", AABB_DTYPE), ("index1", "<u4"), ("index2", "<u4"), ("type1", "u1"), ("type2", "u1"), ("pad", "u1", 2)]
)
assert (RAY_DTYPE.itemsize, RAYCAST_DTYPE.itemsize, AABB_DTYPE.itemsize, NODE_DTYPE.itemsize) == (24, 36, 24, 60)


class Config(C.Structure):
    _fields_ = [("device", C.c_int), ("max_distance", C.c_float), ("morton_bits_per_axis", C.c_int), ("use_cuda_graph", C.c_int)]


class Stats(C.Structure):
    _fields_ = [
        ("n_entities", C.c_uint32), ("n_colliders", C.c_uint32), ("n_levels", C.c_uint32), ("n_internal_nodes", C.c_uint32),
        ("root", C.c_uint32), ("sort_passes", C.c_uint32), ("kernel_launches", C.c_uint64), ("bvh_bytes", C.c_uint64),
        ("full_builds", C.c_uint64), ("refits", C.c_uint64), ("graph_replays", C.c_uint64),
        ("scene_min", C.c_float * 3), ("scene_max", C.c_float * 3), ("key_bits", C.c_uint32), ("fused_keys", C.c_uint32),
    ]


class CommInfo(C.Structure):
    _fields_ = [("n_ranks", C.c_int), ("rank", C.c_int), ("single_process", C.c_int), ("gather_mode", C.c_int), ("nccl_version", C.c_int),
                ("window_bytes", C.c_uint64), ("gathers", C.c_uint64), ("replications", C.c_uint64)]


# every symbol include/b2col.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "b2col_abi_version", "b2col_device_count", "b2col_create", "b2col_destroy", "b2col_last_error", "b2col_sync", "b2col_stream",
    "b2col_set_hierarchy", "b2col_set_trs", "b2col_set_trs_device", "b2col_bind_trs_device", "b2col_update_trs_rows", "b2col_stage_trs_rows", "b2col_commit_trs_rows", "b2col_fetch_world_rows", "b2col_set_world_matrices", "b2col_set_colliders",
    "b2col_update", "b2col_propagate", "b2col_refit", "b2col_raycast_batch", "b2col_raycast_batch_static", "b2col_raycast_device", "b2col_raycast_hits_batch", "b2col_raycast_hits_device", "b2col_expand_hits", "b2col_upload_ref_bvh", "b2col_export_bitreenode",
    "b2col_get_world_matrices", "b2col_get_world_aabbs", "b2col_get_leaf_order", "b2col_get_stats", "b2col_get_trace_counters", "b2col_read_device",
    "b2col_bvh_blob", "b2col_bvh_export_blob", "b2col_bvh_import_blob", "b2col_event_record", "b2col_event_elapsed_ms",
    "b2col_shard_bounds", "b2col_comm_get_unique_id", "b2col_comm_init_rank", "b2col_comm_init_all", "b2col_comm_finalize",
    "b2col_comm_replicate", "b2col_comm_replicate_all", "b2col_comm_replicate_records", "b2col_comm_raycast_gather",
    "b2col_comm_raycast_gather_all", "b2col_mgpu_raycast_batch", "b2col_comm_info",
]


class B2ColError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"b2col status {status}: {message}")
        self.status = status


_lib = None


def load() -> C.CDLL:
    """Load libb2col.so. Raises if it has not been built — there is nothing to fall back to."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} not found: build it with `python -m engine_b200.build` (nvcc, sm_100a). engine_b200 has no CPU path.")
    lib = C.CDLL(LIB_PATH)
    vp, u32, sz = C.c_void_p, C.c_uint32, C.c_size_t
    lib.b2col_abi_version.restype = C.c_int
    lib.b2col_device_count.argtypes = [C.POINTER(C.c_int)]
    lib.b2col_create.argtypes = [C.POINTER(vp), C.POINTER(Config)]
    lib.b2col_destroy.argtypes = [vp]
    lib.b2col_destroy.restype = None
    lib.b2col_last_error.argtypes = [vp]
    lib.b2col_last_error.restype = C.c_char_p
    lib.b2col_sync.argtypes = [vp]
    lib.b2col_stream.argtypes = [vp]
    lib.b2col_stream.restype = vp
    lib.b2col_set_hierarchy.argtypes = [vp, u32, vp]
    lib.b2col_set_trs.argtypes = [vp, u32, vp, vp, vp]
    lib.b2col_set_trs_device.argtypes = [vp, u32, vp, vp, vp]
    lib.b2col_bind_trs_device.argtypes = [vp, u32, vp, vp, vp]
    lib.b2col_set_world_matrices.argtypes = [vp, u32, vp]
    lib.b2col_update_trs_rows.argtypes = [vp, u32, vp, vp, vp, vp]
    lib.b2col_stage_trs_rows.argtypes = [vp, u32, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    lib.b2col_commit_trs_rows.argtypes = [vp, u32]
    lib.b2col_fetch_world_rows.argtypes = [vp, u32, vp, C.POINTER(vp)]
    lib.b2col_set_colliders.argtypes = [vp, u32, vp, vp]
    lib.b2col_update.argtypes = [vp]
    lib.b2col_propagate.argtypes = [vp]
    lib.b2col_refit.argtypes = [vp]
    lib.b2col_raycast_batch.argtypes = [vp, vp, sz, vp, C.c_int]
    lib.b2col_raycast_device.argtypes = [vp, vp, sz, vp, C.c_int]
    lib.b2col_raycast_batch_static.argtypes = [vp, vp, sz, vp]
    lib.b2col_raycast_hits_batch.argtypes = [vp, vp, sz, vp]
    lib.b2col_raycast_hits_device.argtypes = [vp, vp, sz, vp]
    lib.b2col_expand_hits.argtypes = [vp, vp, sz, vp, C.c_int]
    lib.b2col_upload_ref_bvh.argtypes = [vp, vp, sz]
    lib.b2col_export_bitreenode.argtypes = [vp, vp, sz, C.POINTER(sz)]
    lib.b2col_get_world_matrices.argtypes = [vp, u32, u32, vp]
    lib.b2col_get_world_aabbs.argtypes = [vp, vp]
    lib.b2col_get_leaf_order.argtypes = [vp, vp]
    lib.b2col_get_stats.argtypes = [vp, C.POINTER(Stats)]
    lib.b2col_read_device.argtypes = [vp, vp, sz, vp]
    lib.b2col_get_trace_counters.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.b2col_bvh_blob.argtypes = [vp, C.POINTER(vp), C.POINTER(sz)]
    lib.b2col_bvh_import_blob.argtypes = [vp, vp, sz]
    lib.b2col_bvh_export_blob.argtypes = [vp, vp, sz]
    lib.b2col_shard_bounds.argtypes = [sz, C.c_int, C.c_int, C.POINTER(sz), C.POINTER(sz)]
    lib.b2col_shard_bounds.restype = None
    lib.b2col_comm_get_unique_id.argtypes = [vp]
    lib.b2col_comm_init_rank.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.b2col_comm_init_all.argtypes = [C.POINTER(vp), C.c_int]
    lib.b2col_comm_finalize.argtypes = [vp]
    lib.b2col_comm_replicate.argtypes = [vp, C.c_int, u32]
    lib.b2col_comm_replicate_all.argtypes = [C.POINTER(vp), C.c_int, C.c_int]
    lib.b2col_comm_replicate_records.argtypes = [vp, C.c_int, u32]
    lib.b2col_comm_raycast_gather.argtypes = [vp, vp, sz, C.c_int, C.POINTER(vp)]
    lib.b2col_comm_raycast_gather_all.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(vp), sz, C.c_int, C.POINTER(vp)]
    lib.b2col_mgpu_raycast_batch.argtypes = [C.POINTER(vp), C.c_int, vp, sz, vp, C.c_int]
    lib.b2col_comm_info.argtypes = [vp, C.POINTER(CommInfo)]
    lib.b2col_event_record.argtypes = [vp, C.c_int]
    lib.b2col_event_elapsed_ms.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_float)]
    _lib = lib
    return lib


def ptr(a) -> C.c_void_p:
    """Host pointer of a contiguous numpy array, or the address of a torch tensor (host or device)."""
    if a is None:
        return C.c_void_p(0)
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return C.c_void_p(a.ctypes.data)
    if isinstance(a, int):
        return C.c_void_p(a)
    if hasattr(a, "data_ptr"):
        assert a.is_contiguous()
        return C.c_void_p(a.data_ptr())
    raise TypeError(type(a))
