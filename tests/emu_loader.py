"""TEST INFRASTRUCTURE: builds and loads tests/emu/libemu.so (host execution of the kernel bodies)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "emu", "emu.cpp")
LIB = os.path.join(HERE, "emu", "libemu.so")
DEPS = [SRC] + [os.path.join(ROOT, "engine_b200", "csrc", f) for f in ("b2col_math.cuh", "b2col_bvh.cuh")]

NODE64 = np.dtype([("lo0", "<f4", 3), ("hi0", "<f4", 3), ("ref0", "<u4"), ("aux0", "<u4"), ("lo1", "<f4", 3), ("hi1", "<f4", 3), ("ref1", "<u4"), ("aux1", "<u4")])
assert NODE64.itemsize == 64
QNODE = np.dtype([("w", "<u4", 8)])
LEAFREC = np.dtype([("lo", "<f4", 3), ("hi", "<f4", 3), ("entity", "<u4"), ("prim", "<u4")])
GRID = np.dtype([("gmin", "<f4", 3), ("cell", "<f4", 3), ("inv_cell", "<f4", 3)])
assert QNODE.itemsize == 32 and LEAFREC.itemsize == 32 and GRID.itemsize == 36
LEAF_BIT = 0x80000000


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Emu:
    def __init__(self):
        if not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in DEPS):
            subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-std=c++17", "-Wno-unknown-pragmas", "-o", LIB, SRC])
        self.lib = C.CDLL(LIB)
        self.lib.emu_export.restype = C.c_size_t
        self.lib.emu_trace_reftopo.restype = C.c_int

    def propagate(self, parent, pos, rot, scl):
        parent = np.ascontiguousarray(parent, np.uint32)
        n = parent.shape[0] - 1
        world = np.zeros((n + 1, 16), np.float32)
        pos, rot, scl = (np.ascontiguousarray(a, np.float32) for a in (pos, rot, scl))
        self.lib.emu_propagate(C.c_uint32(n), _p(parent), _p(pos), _p(rot), _p(scl), _p(world))
        return world

    def refit(self, entity, local, world):
        entity = np.ascontiguousarray(entity, np.uint32)
        local = np.ascontiguousarray(local, np.float32).reshape(-1, 6)
        world = np.ascontiguousarray(world, np.float32)
        out = np.zeros((entity.shape[0], 6), np.float32)
        self.lib.emu_refit(C.c_uint32(entity.shape[0]), _p(entity), _p(local), _p(world), _p(out))
        return out

    @staticmethod
    def auto_bits(n):
        """the library's automatic key width (b2col_update): (bits per axis, dropped low bits)"""
        key_bits = 32 if n <= (1 << 21) else (40 if n <= (1 << 25) else 48)
        bits = (key_bits + 2) // 3
        return bits, 3 * bits - key_bits

    def build(self, world_aabb, entity, bits=0, order=None, T=1000.0, key_bits=0, plan_bounds=None):
        """bits: bits per axis (cfg.morton_bits_per_axis); key_bits: total key width, as b2col_update's key-width policy picks it; neither: the static width"""
        world_aabb = np.ascontiguousarray(world_aabb, np.float32).reshape(-1, 6)
        entity = np.ascontiguousarray(entity, np.uint32)
        n = entity.shape[0]
        if key_bits:
            bits = (key_bits + 2) // 3
            drop = 3 * bits - key_bits
        else:
            bits, drop = (bits, 0) if bits else self.auto_bits(n)
        nodes = np.zeros(max(n - 1, 1), NODE64)
        qnodes = np.zeros(max(n - 1, 1), QNODE)
        leaves = np.zeros(n, LEAFREC)
        grid = np.zeros(1, GRID)
        sorted_prim = np.zeros(n, np.uint32)
        sorted_keys = np.zeros(n, np.uint64)
        root = C.c_uint32(0)
        mismatch = C.c_int(0)
        has_wild = C.c_uint32(0)
        if order is not None:
            order = np.ascontiguousarray(order, np.uint32)
        if plan_bounds is not None:  # (min.xyz, max.xyz) of the Morton plan, as b2col_get_stats reports them (scene_min / scene_max)
            plan_bounds = np.ascontiguousarray(np.concatenate([np.asarray(plan_bounds[0], np.float32), np.asarray(plan_bounds[1], np.float32)]), np.float32)
        self.lib.emu_build(C.c_uint32(n), _p(world_aabb), _p(entity), C.c_int(bits), C.c_int(drop), C.c_float(T), _p(nodes), _p(qnodes), _p(leaves), _p(grid),
                           _p(sorted_prim), _p(sorted_keys), C.byref(root), _p(order), C.byref(mismatch), C.byref(has_wild), _p(plan_bounds))
        assert mismatch.value == 0, f"grid-space builder disagrees with the exact builder (flags {mismatch.value})"
        return dict(n=n, nodes=nodes, qnodes=qnodes, leaves=leaves, grid=grid, leaf_entity=leaves["entity"].copy(), sorted_prim=sorted_prim,
                    sorted_keys=sorted_keys, root=root.value, world_aabb=world_aabb, has_wild=int(has_wild.value))

    def morton(self, centroids, cmin, cmax, key_bits):
        """(plan dict, keys) of engine_b200/csrc/b2col_math.cuh make_morton_plan / morton_key"""
        c = np.ascontiguousarray(centroids, np.float32).reshape(-1, 3)
        cmin, cmax = np.ascontiguousarray(cmin, np.float32), np.ascontiguousarray(cmax, np.float32)
        plan = np.zeros(7, np.int32)
        keys = np.zeros(c.shape[0], np.uint64)
        self.lib.emu_morton(C.c_uint32(c.shape[0]), _p(c), _p(cmin), _p(cmax), C.c_int(key_bits), _p(plan), _p(keys))
        names = ("a", "b", "c", "shift_a", "nc", "d2", "drop")
        return dict(zip(names, (int(v) for v in plan))), keys

    def trace_lbvh(self, tree, rays, T=1000.0):
        from oracle.oracle import RAYCAST_DTYPE

        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], RAYCAST_DTYPE)
        self.lib.emu_trace_lbvh(_p(tree["qnodes"]), _p(tree["leaves"]), _p(tree["grid"]), C.c_uint32(tree["root"]), C.c_uint32(tree["n"]), _p(rays),
                                C.c_size_t(rays.shape[0]), _p(out), C.c_float(T), C.c_uint32(tree.get("has_wild", 0)))
        return out

    def trace_reftopo(self, nodes60, rays, T=1000.0):
        from oracle.oracle import NODE_DTYPE, RAYCAST_DTYPE

        nodes60 = np.ascontiguousarray(nodes60, NODE_DTYPE)
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], RAYCAST_DTYPE)
        ov = self.lib.emu_trace_reftopo(_p(nodes60), C.c_uint32(nodes60.shape[0]), _p(rays), C.c_size_t(rays.shape[0]), _p(out), C.c_float(T))
        assert ov == 0, "stack overflow"
        return out

    def export(self, tree):
        from oracle.oracle import NODE_DTYPE

        n = tree["n"]
        out = np.zeros(2 * max(n, 1), NODE_DTYPE)
        k = self.lib.emu_export(C.c_uint32(n), _p(tree["nodes"]), C.c_uint32(tree["root"]), _p(tree["leaves"]), _p(out), C.c_size_t(out.shape[0]))
        return out[:k].copy()
