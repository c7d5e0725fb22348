"""ctypes binding of include/b2col_gltf.h (libb2col_gltf.so, host-only C++): collider ingestion from .glb files.

``load_colliders`` returns what the reference's loadGLTF (reference src/gltf_loader.cpp:101-684) would create on
the collision path — entities in creation order with parent + TRS, and the colliders — as arrays that plug into
``CollisionSystem.set_hierarchy / set_trs / set_colliders``.  ``append_to_scene`` adds a loaded file to an
``engine_b200.synth.Scene``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb2col_gltf.so")
OK, ERR_IO, ERR_FORMAT, ERR_UNSUPPORTED = range(4)

SYMBOLS = [
    "b2col_gltf_load_file", "b2col_gltf_load_memory", "b2col_gltf_free", "b2col_gltf_last_error", "b2col_gltf_entity_count",
    "b2col_gltf_collider_count", "b2col_gltf_get_entities", "b2col_gltf_get_colliders", "b2col_gltf_entity_name",
]


class GltfError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"b2col_gltf status {status}: {message}")
        self.status = status


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not found: build it with `python -m engine_b200.build`")
        lib = C.CDLL(LIB_PATH)
        vp, u32 = C.c_void_p, C.c_uint32
        lib.b2col_gltf_load_file.argtypes = [C.c_char_p, u32, u32, C.c_int, C.POINTER(vp)]
        lib.b2col_gltf_load_memory.argtypes = [vp, C.c_size_t, C.c_char_p, u32, u32, C.c_int, C.POINTER(vp)]
        lib.b2col_gltf_free.argtypes = [vp]
        lib.b2col_gltf_free.restype = None
        lib.b2col_gltf_last_error.restype = C.c_char_p
        lib.b2col_gltf_entity_count.argtypes = [vp]
        lib.b2col_gltf_entity_count.restype = u32
        lib.b2col_gltf_collider_count.argtypes = [vp]
        lib.b2col_gltf_collider_count.restype = u32
        lib.b2col_gltf_get_entities.argtypes = [vp, vp, vp, vp, vp]
        lib.b2col_gltf_get_entities.restype = None
        lib.b2col_gltf_get_colliders.argtypes = [vp, vp, vp]
        lib.b2col_gltf_get_colliders.restype = None
        lib.b2col_gltf_entity_name.argtypes = [vp, u32]
        lib.b2col_gltf_entity_name.restype = C.c_char_p
        _lib = lib
    return _lib


def load_colliders(source, first_entity: int = 1, root_parent: int = 0, strict: bool = True, stem: str = "gltf") -> dict:
    """source: a path, or the bytes of a .glb.  Keys: ids, parent, pos, rot (xyzw), scl, tags, collider_entity, collider_aabb."""
    lib = load()
    h = C.c_void_p()
    if isinstance(source, (bytes, bytearray, memoryview)):
        buf = bytes(source)
        st = lib.b2col_gltf_load_memory(buf, len(buf), stem.encode(), first_entity, root_parent, int(strict), C.byref(h))
    else:
        st = lib.b2col_gltf_load_file(os.fspath(source).encode(), first_entity, root_parent, int(strict), C.byref(h))
    if st != OK:
        raise GltfError(st, (lib.b2col_gltf_last_error() or b"").decode())
    try:
        n, k = lib.b2col_gltf_entity_count(h), lib.b2col_gltf_collider_count(h)
        parent = np.zeros(n, np.uint32)
        pos, rot, scl = np.zeros((n, 3), np.float32), np.zeros((n, 4), np.float32), np.zeros((n, 3), np.float32)
        ent, box = np.zeros(k, np.uint32), np.zeros((k, 6), np.float32)
        lib.b2col_gltf_get_entities(h, parent.ctypes.data, pos.ctypes.data, rot.ctypes.data, scl.ctypes.data)
        lib.b2col_gltf_get_colliders(h, ent.ctypes.data, box.ctypes.data)
        tags = [(lib.b2col_gltf_entity_name(h, i) or b"").decode() for i in range(n)]
    finally:
        lib.b2col_gltf_free(h)
    return dict(ids=np.arange(first_entity, first_entity + n, dtype=np.uint32), parent=parent, pos=pos, rot=rot, scl=scl, tags=tags,
                collider_entity=ent, collider_aabb=box)


def append_to_scene(scene, source, root_parent: int = 0, strict: bool = True, stem: str = "gltf") -> int:
    """Load a .glb behind the entities already in ``scene`` (engine_b200.synth.Scene; arrays indexed by entity ID,
    row 0 unused).  Returns the file-root entity's ID, as loadGLTF does."""
    first = scene.n_entities + 1
    g = load_colliders(source, first_entity=first, root_parent=root_parent, strict=strict, stem=stem)
    scene.parent = np.concatenate([scene.parent, g["parent"]]).astype(np.uint32)
    scene.pos = np.concatenate([scene.pos, g["pos"]]).astype(np.float32)
    scene.rot = np.concatenate([scene.rot, g["rot"]]).astype(np.float32)
    scene.scl = np.concatenate([scene.scl, g["scl"]]).astype(np.float32)
    scene.collider_entity = np.concatenate([scene.collider_entity, g["collider_entity"]]).astype(np.uint32)
    scene.collider_aabb = np.concatenate([scene.collider_aabb.reshape(-1, 6), g["collider_aabb"]]).astype(np.float32)
    return first
