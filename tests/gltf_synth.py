"""Synthetic .glb files for the collider-ingestion tests: seeded node hierarchies with TRS and
`matrix` nodes, unnamed nodes, meshes with 1..3 primitives (each a real triangle with POSITION / NORMAL /
TEXCOORD_0 + indices, so that the UNMODIFIED reference loader — which builds vertex data and tangents — accepts
them), primitives without POSITION, several scenes with a non-zero default."""
from __future__ import annotations

import json
import struct

import numpy as np


def _quat(rng):
    q = rng.standard_normal(4)
    return (q / np.linalg.norm(q)).tolist()


def _trs_matrix(t, q, s):
    x, y, z, w = q
    r = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    m = np.eye(4)
    m[:3, :3] = r * np.array(s)[None, :]
    m[:3, 3] = t
    return m.T.reshape(-1).tolist()  # column-major


def make_glb(seed: int = 0, n_nodes: int = 12, n_meshes: int = 4, flavour: str = "mixed") -> bytes:
    rng = np.random.default_rng(seed)
    blob = bytearray()
    views, accessors, meshes = [], [], []

    def add_view(data: bytes, target=None):
        while len(blob) % 4:
            blob.append(0)
        v = {"buffer": 0, "byteOffset": len(blob), "byteLength": len(data)}
        if target:
            v["target"] = target
        blob.extend(data)
        views.append(v)
        return len(views) - 1

    def add_accessor(arr: np.ndarray, ctype: int, typ: str, with_minmax=False, target=None):
        view = add_view(arr.tobytes(), target)
        a = {"bufferView": view, "componentType": ctype, "count": int(arr.shape[0]), "type": typ}
        if with_minmax:
            a["min"] = [float(x) for x in arr.min(0)]
            a["max"] = [float(x) for x in arr.max(0)]
        accessors.append(a)
        return len(accessors) - 1

    for mi in range(n_meshes):
        prims = []
        for _ in range(int(rng.integers(1, 4))):
            c = rng.uniform(-3, 3, 3)
            pos = (c + rng.uniform(-1.5, 1.5, (3, 3))).astype(np.float32)
            nrm = np.tile(np.array([[0, 0, 1]], np.float32), (3, 1))
            uv = rng.uniform(0, 1, (3, 2)).astype(np.float32)
            attrs = {"NORMAL": add_accessor(nrm, 5126, "VEC3", target=34962), "TEXCOORD_0": add_accessor(uv, 5126, "VEC2", target=34962)}
            if not (flavour == "mixed" and rng.random() < 0.15):  # some primitives have no POSITION: the reference skips them
                attrs["POSITION"] = add_accessor(pos, 5126, "VEC3", with_minmax=True, target=34962)
            prims.append({"attributes": attrs, "indices": add_accessor(np.array([0, 1, 2], np.uint16), 5123, "SCALAR", target=34963), "mode": 4})
        meshes.append({"name": f"mesh{mi}", "primitives": prims})

    nodes = []
    for ni in range(n_nodes):
        node = {}
        if rng.random() < 0.8:
            node["name"] = f"node{ni}"
        kind = rng.random()
        t, q, s = rng.uniform(-10, 10, 3).tolist(), _quat(rng), rng.uniform(0.3, 3.0, 3).tolist()
        if kind < 0.35:
            node["matrix"] = _trs_matrix(t, q, s)
        elif kind < 0.9:
            if rng.random() < 0.8:
                node["translation"] = t
            if rng.random() < 0.8:
                node["rotation"] = q
            if rng.random() < 0.8:
                node["scale"] = s
        if rng.random() < 0.7 and n_meshes:
            node["mesh"] = int(rng.integers(0, n_meshes))
        nodes.append(node)
    # a forest: node i > 0 hangs under an earlier node or is a root
    roots = [0]
    for ni in range(1, n_nodes):
        if rng.random() < 0.2:
            roots.append(ni)
        else:
            p = int(rng.integers(0, ni))
            nodes[p].setdefault("children", []).append(ni)
    split = max(1, len(roots) // 2)
    doc = {
        "asset": {"version": "2.0", "generator": "engine_b200 tests"},
        "scene": 1,
        "scenes": [{"name": "unused", "nodes": roots[split:] or roots[:1]}, {"name": "main", "nodes": roots[:split]}],
        "nodes": nodes,
        "meshes": meshes,
        "accessors": accessors,
        "bufferViews": views,
        "buffers": [{"byteLength": len(blob)}],
    }
    if flavour == "single_scene":
        del doc["scene"]
        doc["scenes"] = [{"nodes": roots}]
    js = json.dumps(doc, separators=(",", ":")).encode()
    js += b" " * (-len(js) % 4)
    while len(blob) % 4:
        blob.append(0)
    total = 12 + 8 + len(js) + 8 + len(blob)
    return b"glTF" + struct.pack("<II", 2, total) + struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(blob), 0x004E4942) + bytes(blob)
