"""TEST INFRASTRUCTURE — ctypes loaders for the two CPU checkers.

* :class:`Port`  — ``oracle/liboracle_port.so`` (oracle/port.c, the plain-C restatement).
* :class:`RefEngine` — ``oracle/_ref/libref_engine.so``: the UNMODIFIED reference collision path
  compiled from ``/root/reference`` (oracle/Makefile).  Present wherever the build container's
  snapshot travels; ``RefEngine.available()`` says whether it is.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import this module.  ``engine_b200`` never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle_port.so")
REF_SO = os.path.join(HERE, "_ref", "libref_engine.so")
# reference Scene/ECS/TransformSystem + engine_b200/shim's replacement CollisionSystem (the drop-in test)
DROPIN_SO = os.path.join(HERE, "_ref", "libdropin_engine.so")

# Layouts of the reference's PODs (include/system_collisions.h:13-24,40-48; component_collider.h:7-10)
RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("direction", "<f4", 3)])
RAYCAST_DTYPE = np.dtype(
    [("location", "<f4", 3), ("normal", "<f4", 3), ("hit_entity", "<u4"), ("distance", "<f4"), ("hit", "u1"), ("pad", "u1", 3)]
)
AABB_DTYPE = np.dtype([("min", "<f4", 3), ("max", "<f4", 3)])
NODE_DTYPE = np.dtype(
    [("box1", AABB_DTYPE), ("box2", AABB_DTYPE), ("index1", "<u4"), ("index2", "<u4"), ("type1", "u1"), ("type2", "u1"), ("pad", "u1", 2)]
)
assert RAY_DTYPE.itemsize == 24 and RAYCAST_DTYPE.itemsize == 36 and AABB_DTYPE.itemsize == 24 and NODE_DTYPE.itemsize == 60

TYPE_BV, TYPE_ENTITY, TYPE_EMPTY = 0, 1, 2


def build(force: bool = False) -> None:
    """(Re)build the checkers with oracle/Makefile. Building the checker is not using it."""
    if force or not os.path.exists(PORT_SO) or (os.path.isdir("/root/reference") and not os.path.exists(REF_SO)):
        subprocess.check_call(["make", "-C", HERE, "all"], stdout=subprocess.DEVNULL)


def _p(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None:
        a = a.reshape(shape)
    return a


def as_rays(rays) -> np.ndarray:
    """Accept (n,6) float32 or RAY_DTYPE; return contiguous RAY_DTYPE view."""
    rays = np.ascontiguousarray(rays)
    if rays.dtype != RAY_DTYPE:
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6).view(RAY_DTYPE).reshape(-1)
    return rays


def as_aabbs(b) -> np.ndarray:
    b = np.ascontiguousarray(b)
    if b.dtype != AABB_DTYPE:
        b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1, 6).view(AABB_DTYPE).reshape(-1)
    return b


class Port:
    """oracle/port.c"""

    def __init__(self):
        build()
        self.lib = C.CDLL(PORT_SO)
        L = self.lib
        L.orc_build_binned_sah.restype = C.c_size_t
        L.orc_tree_leaves_dfs.restype = C.c_size_t
        L.orc_ray_box.restype = C.c_int

    # -- transform ---------------------------------------------------------------------------
    def transform_update(self, parent, pos, rot_xyzw, scl, world_prev=None) -> np.ndarray:
        """Arrays indexed by entity ID (row 0 unused). Returns (n+1,16) column-major matrices."""
        parent = np.ascontiguousarray(parent, dtype=np.uint32)
        n = parent.shape[0] - 1
        pos, rot, scl = _f32(pos, (n + 1, 3)), _f32(rot_xyzw, (n + 1, 4)), _f32(scl, (n + 1, 3))
        world = np.zeros((n + 1, 16), np.float32) if world_prev is None else np.array(world_prev, np.float32, copy=True)
        self.lib.orc_transform_update(C.c_size_t(n), _p(parent), _p(pos), _p(rot), _p(scl), _p(world))
        return world

    def refit(self, entity, local_aabb, world16) -> np.ndarray:
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        local = as_aabbs(local_aabb)
        world16 = _f32(world16)
        out = np.zeros(entity.shape[0], AABB_DTYPE)
        self.lib.orc_refit(C.c_size_t(entity.shape[0]), _p(entity), _p(local), _p(world16), _p(out))
        return out

    def transform_box(self, aabb6, mat16) -> np.ndarray:
        a, m = _f32(aabb6, (6,)), _f32(mat16, (16,))
        out = np.zeros(6, np.float32)
        self.lib.orc_transform_box(_p(a), _p(m), _p(out))
        return out

    # -- queries -----------------------------------------------------------------------------
    def ray_box(self, ray6, aabb6, T=1000.0):
        r, a = _f32(ray6, (6,)), _f32(aabb6, (6,))
        t, side = C.c_float(), C.c_int()
        hit = self.lib.orc_ray_box(_p(r), _p(a), C.c_float(T), C.byref(t), C.byref(side))
        return bool(hit), np.float32(t.value), side.value

    def normalize(self, v3) -> np.ndarray:
        v = _f32(v3, (3,))
        out = np.zeros(3, np.float32)
        self.lib.orc_normalize(_p(v), _p(out))
        return out

    def raycast_tree(self, nodes, rays, nthreads=1) -> np.ndarray:
        nodes = np.ascontiguousarray(nodes, dtype=NODE_DTYPE)
        rays = as_rays(rays)
        out = np.zeros(rays.shape[0], RAYCAST_DTYPE)
        self.lib.orc_raycast_batch(_p(nodes), C.c_size_t(nodes.shape[0]), _p(rays), C.c_size_t(rays.shape[0]), _p(out), C.c_int(nthreads))
        return out

    def raycast_brute(self, entity, boxes, rays, nthreads=1, want_ties=False):
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        boxes = as_aabbs(boxes)
        rays = as_rays(rays)
        out = np.zeros(rays.shape[0], RAYCAST_DTYPE)
        ties = np.zeros(rays.shape[0], np.uint32)
        self.lib.orc_brute_batch(
            _p(entity), _p(boxes), C.c_size_t(entity.shape[0]), _p(rays), C.c_size_t(rays.shape[0]), _p(out), _p(ties), C.c_int(nthreads)
        )
        return (out, ties) if want_ties else out

    # -- trees -------------------------------------------------------------------------------
    def build_binned_sah(self, entity, boxes) -> np.ndarray:
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        boxes = as_aabbs(boxes)
        n = entity.shape[0]
        nodes = np.zeros(max(2 * n - 1, 0), NODE_DTYPE)
        k = self.lib.orc_build_binned_sah(_p(entity), _p(boxes), C.c_size_t(n), _p(nodes))
        return nodes[:k].copy()

    def tree_leaves_dfs(self, nodes):
        nodes = np.ascontiguousarray(nodes, dtype=NODE_DTYPE)
        cap = 2 * nodes.shape[0]
        ent = np.zeros(cap, np.uint32)
        box = np.zeros(cap, AABB_DTYPE)
        k = self.lib.orc_tree_leaves_dfs(_p(nodes), C.c_size_t(nodes.shape[0]), _p(ent), _p(box))
        return ent[:k].copy(), box[:k].copy()


class RefEngine:
    """The real reference (engine::Scene + TransformSystem + CollisionSystem) behind oracle/ref_harness.cpp."""

    @staticmethod
    def available() -> bool:
        if not os.path.exists(REF_SO) and os.path.isdir("/root/reference"):
            build()
        return os.path.exists(REF_SO)

    SO = REF_SO

    def __init__(self):
        if not self.available() or not os.path.exists(self.SO):
            raise RuntimeError(f"{self.SO} not built (needs /root/reference)")
        self.lib = C.CDLL(self.SO)
        L = self.lib
        for name, rt in (("refh_create", C.c_void_p), ("refh_sizeof", C.c_size_t), ("refh_create_entities", C.c_uint32),
                         ("refh_scene_update", C.c_double), ("refh_transform_update", C.c_double), ("refh_refit", C.c_size_t),
                         ("refh_bvh_size", C.c_size_t), ("refh_raycast", C.c_double), ("refh_ray_box", C.c_int)):
            try:
                getattr(L, name).restype = rt
            except AttributeError:  # the drop-in build has no access to the reference's file-static helpers
                pass
        self.h = C.c_void_p(L.refh_create())
        self.n_entities = 0

    def close(self):
        if self.h:
            self.lib.refh_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def hw_threads(self) -> int:
        return int(self.lib.refh_hw_threads())

    def create_entities(self, parent, pos, rot_xyzw, scl) -> np.ndarray:
        """Rows are NEW entities in creation order (no dummy row 0). Returns their IDs."""
        parent = np.ascontiguousarray(parent, dtype=np.uint32)
        n = parent.shape[0]
        pos, rot, scl = _f32(pos, (n, 3)), _f32(rot_xyzw, (n, 4)), _f32(scl, (n, 3))
        first = self.lib.refh_create_entities(self.h, C.c_size_t(n), _p(parent), _p(pos), _p(rot), _p(scl))
        self.n_entities += n
        return np.arange(first, first + n, dtype=np.uint32)

    def add_colliders(self, entity, local_aabb) -> None:
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        local = as_aabbs(local_aabb)
        self.lib.refh_add_colliders(self.h, C.c_size_t(entity.shape[0]), _p(entity), _p(local))

    def set_trs(self, entity, pos, rot_xyzw, scl) -> None:
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        n = entity.shape[0]
        pos, rot, scl = _f32(pos, (n, 3)), _f32(rot_xyzw, (n, 4)), _f32(scl, (n, 3))
        self.lib.refh_set_trs(self.h, C.c_size_t(n), _p(entity), _p(pos), _p(rot), _p(scl))

    def set_static(self, entity, is_static: bool = True) -> None:
        """TransformComponent::is_static of the given entities (reference include/component_transform.h:24)"""
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        self.lib.refh_set_static(self.h, C.c_size_t(entity.shape[0]), _p(entity), C.c_int(1 if is_static else 0))

    SHIM_OPTIONS = {"collider_check_interval": 0, "rebuild_interval": 1, "skip_static_rows": 2, "write_back_world_matrices": 3, "static_rescan_interval": 4,
                    "use_devices": 5, "multi_gpu_min_rays": 6, "full_uploads": 10, "sparse_uploads": 11, "devices_in_use": 12}

    def shim_option(self, name: str, value: int = 0) -> int:
        """knobs / counters of the drop-in shims (no-op in the reference build)"""
        self.lib.refh_shim_option.restype = C.c_long
        return int(self.lib.refh_shim_option(self.h, C.c_int(self.SHIM_OPTIONS[name]), C.c_long(value)))

    def scene_update(self) -> float:
        return float(self.lib.refh_scene_update(self.h))

    def transform_update(self) -> float:
        return float(self.lib.refh_transform_update(self.h))

    def refit(self, n_colliders: int):
        ent = np.zeros(n_colliders, np.uint32)
        box = np.zeros(n_colliders, AABB_DTYPE)
        secs = C.c_double()
        k = self.lib.refh_refit(self.h, _p(ent), _p(box), C.byref(secs))
        assert k == n_colliders, (k, n_colliders)
        return ent, box, secs.value

    def world_matrices(self, entity) -> np.ndarray:
        entity = np.ascontiguousarray(entity, dtype=np.uint32)
        out = np.zeros((entity.shape[0], 16), np.float32)
        self.lib.refh_get_world_matrices(self.h, C.c_size_t(entity.shape[0]), _p(entity), _p(out))
        return out

    def bvh(self) -> np.ndarray:
        n = self.lib.refh_bvh_size(self.h)
        out = np.zeros(n, NODE_DTYPE)
        self.lib.refh_bvh_get(self.h, _p(out))
        out["pad"] = 0  # the two tail padding bytes are indeterminate in the reference
        return out

    def set_bvh(self, nodes) -> None:
        nodes = np.ascontiguousarray(nodes, dtype=NODE_DTYPE)
        self.lib.refh_bvh_set(self.h, _p(nodes), C.c_size_t(nodes.shape[0]))

    def raycast(self, rays, nthreads=1, want_secs=False):
        rays = as_rays(rays)
        out = np.zeros(rays.shape[0], RAYCAST_DTYPE)
        secs = self.lib.refh_raycast(self.h, C.c_size_t(rays.shape[0]), _p(rays), _p(out), C.c_int(nthreads))
        return (out, float(secs)) if want_secs else out

    def ray_box(self, ray6, aabb6, T=1000.0):
        r, a = _f32(ray6, (6,)), _f32(aabb6, (6,))
        t, side = C.c_float(), C.c_int()
        hit = self.lib.refh_ray_box(_p(r), _p(a), C.c_float(T), C.byref(t), C.byref(side))
        return bool(hit), np.float32(t.value), side.value

    def transform_box(self, aabb6, mat16) -> np.ndarray:
        a, m = _f32(aabb6, (6,)), _f32(mat16, (16,))
        out = np.zeros(6, np.float32)
        self.lib.refh_transform_box(_p(a), _p(m), _p(out))
        return out

    def normalize(self, v3) -> np.ndarray:
        v = _f32(v3, (3,))
        out = np.zeros(3, np.float32)
        self.lib.refh_normalize(_p(v), _p(out))
        return out

    # -- the reference glTF loader, headless (oracle/gltf_harness.cpp) -------------------------
    def load_gltf(self, path: str, is_static: bool = True, impl: str = "reference") -> dict:
        """loadGLTF (reference src/gltf_loader.cpp:101-684) into this engine's Scene.  Returns what it created on the
        collision path: the new entities in creation order (ids, parent, pos, rot xyzw, scl, tags) and the colliders
        among them (entity, aabb)."""
        L = self.lib
        L.gltfh_load.restype = C.c_uint32
        L.gltfh_entity_count.restype = C.c_uint32
        L.gltfh_dump_colliders.restype = C.c_uint32
        L.gltfh_last_error.restype = C.c_char_p
        before = int(L.gltfh_entity_count(self.h))
        # impl "shim": engine_b200/shim's loadGLTFColliders creating entities in the same reference Scene
        L.gltfh_load_b2.restype = C.c_uint32
        fn = L.gltfh_load if impl == "reference" else L.gltfh_load_b2
        root = int(fn(self.h, path.encode(), C.c_int(1 if is_static else 0)))
        if root == 0:
            raise RuntimeError(f"{impl} glTF load failed: " + (L.gltfh_last_error() or b"").decode())
        n = int(L.gltfh_entity_count(self.h)) - before
        self.n_entities += n
        parent = np.zeros(n, np.uint32)
        pos, rot, scl = np.zeros((n, 3), np.float32), np.zeros((n, 4), np.float32), np.zeros((n, 3), np.float32)
        static = np.zeros(n, np.uint8)
        tags = np.zeros((n, 64), np.uint8)
        L.gltfh_dump_entities(self.h, C.c_uint32(before + 1), C.c_uint32(n), _p(parent), _p(pos), _p(rot), _p(scl), _p(static), _p(tags))
        k = int(L.gltfh_dump_colliders(self.h, None, None))
        ent, box = np.zeros(k, np.uint32), np.zeros((k, 6), np.float32)
        L.gltfh_dump_colliders(self.h, _p(ent), _p(box))
        keep = ent > before
        return dict(root=root, first=before + 1, ids=np.arange(before + 1, before + 1 + n, dtype=np.uint32), parent=parent, pos=pos, rot=rot, scl=scl,
                    is_static=static, tags=[bytes(t).split(b"\0")[0].decode() for t in tags], collider_entity=ent[keep], collider_aabb=box[keep])

    # -- the test game's player controller, headless (oracle/game_harness.cpp) --------------
    GAME_STATE = ("px", "py", "pz", "vx", "vy", "vz", "qx", "qy", "qz", "qw", "yaw", "pitch", "grounded")

    def add_camera(self, pos3, batched: bool = False) -> int:
        """Camera entity + CameraControllerComponent + the controller system at index 0 (reference
        test/src/game.cpp:92-109): the UNMODIFIED reference controller, or (drop-in builds) the batched port."""
        self.lib.gameh_add_camera.restype = C.c_uint32
        cam = int(self.lib.gameh_add_camera(self.h, _p(_f32(pos3, (3,))), C.c_int(1 if batched else 0)))
        if cam == 0:
            raise RuntimeError("the batched controller only exists in the drop-in builds")
        self.n_entities += 1
        return cam

    def run_game(self, input6, dt: float = 1.0 / 60.0) -> np.ndarray:
        """One Scene::Update(dt) per input row (movex, movey, lookx, looky, sprint, jump); returns (frames, 16)
        float32 player state per frame, columns GAME_STATE (+ 3 spare)."""
        inp = _f32(input6).reshape(-1, 6)
        out = np.zeros((inp.shape[0], 16), np.float32)
        self.lib.gameh_run(self.h, C.c_size_t(inp.shape[0]), _p(inp), C.c_float(dt), _p(out))
        return out


DROPIN2_SO = os.path.join(os.path.dirname(DROPIN_SO), "libdropin2_engine.so")


class DropinEngine(RefEngine):
    """The reference's Scene / ECS / TransformSystem with engine_b200/shim's CollisionSystem compiled in
    place of the reference's (oracle/_ref/libdropin_engine.so, needs a GPU at run time)."""

    SO = DROPIN_SO

    @staticmethod
    def available() -> bool:
        return os.path.exists(DROPIN_SO)

    def raycast_batch(self, rays) -> np.ndarray:
        rays = as_rays(rays)
        out = np.zeros(rays.shape[0], RAYCAST_DTYPE)
        self.lib.refh_raycast_batch(self.h, C.c_size_t(rays.shape[0]), _p(rays), _p(out))
        return out

    def raycast_hits_expanded(self, rays, threads: int = 2):
        """drop-in GetRaycastHits (12-byte results) + ExpandHits on the host: (hits as (n,3) uint32 words, Raycasts)"""
        rays = as_rays(rays)
        hits = np.zeros((rays.shape[0], 3), np.uint32)
        out = np.zeros(rays.shape[0], RAYCAST_DTYPE)
        self.lib.refh_raycast_hits_expanded(self.h, C.c_size_t(rays.shape[0]), _p(rays), _p(hits), _p(out), C.c_int(threads))
        return hits, out

    def batched_counters(self) -> dict:
        out = np.zeros(4, np.uint64)
        self.lib.gameh_batched_counters(self.h, _p(out))
        return dict(zip(("raycast_calls", "rays_cast", "frames", "recasts"), (int(x) for x in out)))

    def set_rebuild_interval(self, interval: int) -> None:
        """engine_b200/shim CollisionSystem::rebuild_interval: refit-only frames between full rebuilds."""
        self.lib.gameh_set_rebuild_interval(self.h, C.c_uint(interval))


class Dropin2Engine(DropinEngine):
    """As DropinEngine, with engine_b200/shim/with_transform's TransformSystem replacing the reference's too
    (oracle/_ref/libdropin2_engine.so): TRS gather -> GPU propagation -> world_matrix scatter, matrices kept on
    the device for the collision refit."""

    SO = DROPIN2_SO

    @staticmethod
    def available() -> bool:
        return os.path.exists(DROPIN2_SO)
