"""Collider ingestion from .glb files (include/b2col_gltf.h, engine_b200/host/gltf_colliders.cpp)
against the UNMODIFIED reference loader (reference src/gltf_loader.cpp run headless by oracle/gltf_harness.cpp):
through tests/golden/gltf.npz (made by tests/golden/make_golden.py), and live where oracle/_ref is built.
Host-only code: these tests need no GPU."""
import ctypes as C
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN
from engine_b200 import gltf
from gltf_synth import make_glb
from oracle.oracle import RefEngine

KEYS = ("parent", "pos", "rot", "scl", "collider_entity", "collider_aabb")


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def same(got, want, what):
    assert len(got["parent"]) == len(want["parent"]), what
    for k in KEYS:
        assert (bits(got[k]) == bits(want[k])).all(), f"{what}: {k} differs"
    assert list(got["tags"]) == list(want["tags"]), what


@pytest.fixture(scope="module")
def golden():
    z = np.load(os.path.join(GOLDEN, "gltf.npz"))
    return z, bytes(z["names"]).decode().split("\n")


def test_library_exports_every_declared_symbol():
    lib = gltf.load()
    header = open(os.path.join(os.path.dirname(GOLDEN), "..", "include", "b2col_gltf.h")).read()
    for sym in gltf.SYMBOLS:
        assert hasattr(lib, sym), sym
        assert sym + "(" in header, f"{sym} not declared in include/b2col_gltf.h"


def test_synthetic_files_match_reference_golden(golden):
    z, names = golden
    n = 0
    for key in names:
        if not key.startswith("synth"):
            continue
        data = bytes(z[key + "/glb"])
        seed = int(key[5:])
        assert data == make_glb(seed, n_nodes=6 + 3 * seed, n_meshes=1 + seed % 5, flavour=["mixed", "single_scene"][seed % 2]), "generator drifted from the fixture"
        got = gltf.load_colliders(data, stem=key)
        want = {k: z[f"{key}/{k}"] for k in KEYS}
        want["tags"] = bytes(z[key + "/tags"]).decode().split("\n")
        same(got, want, key)
        n += 1
    assert n == 12


@pytest.mark.skipif(not os.path.isdir("/root/reference/test/res/models"), reason="reference assets not present on this machine")
def test_reference_models_match_reference_golden(golden):
    z, names = golden
    for key in names:
        if not key.startswith("model_"):
            continue
        path = f"/root/reference/test/res/models/{key[6:]}.glb"
        if key + "/error" in z.files:  # the reference loader throws on this file; so must strict mode, with the same message
            with pytest.raises(gltf.GltfError) as e:
                gltf.load_colliders(path)
            assert e.value.status == gltf.ERR_UNSUPPORTED
            assert str(e.value).split(": ", 1)[1] in bytes(z[key + "/error"]).decode()
            assert len(gltf.load_colliders(path, strict=False)["collider_entity"]) >= 1
            continue
        want = {k: z[f"{key}/{k}"] for k in KEYS}
        want["tags"] = bytes(z[key + "/tags"]).decode().split("\n")
        same(gltf.load_colliders(path), want, key)


@pytest.mark.skipif(not RefEngine.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_live_reference_loader_and_scene_facing_shim(tmp_path):
    """Fresh seeds through the live reference loader, the C ABI, and engine_b200/shim's loadGLTFColliders creating
    entities inside the reference's own Scene."""
    for seed in range(100, 116):
        data = make_glb(seed, n_nodes=4 + (seed % 9) * 5, n_meshes=1 + seed % 4, flavour=["mixed", "single_scene"][seed % 2])
        path = str(tmp_path / f"s{seed}.glb")
        open(path, "wb").write(data)
        want = RefEngine().load_gltf(path)
        same(gltf.load_colliders(path), want, f"seed {seed} C ABI")
        shim = RefEngine().load_gltf(path, impl="shim")
        same(shim, want, f"seed {seed} loadGLTFColliders in Scene")
        assert (shim["is_static"] == want["is_static"]).all()
    # a second file loaded behind the first one: IDs continue, parents point into the second file
    e1, e2 = RefEngine(), RefEngine()
    a, b = e1.load_gltf(path), e1.load_gltf(path)
    e2.load_gltf(path, impl="shim")
    b2 = e2.load_gltf(path, impl="shim")
    same(b2, b, "second load")
    g = gltf.load_colliders(path, first_entity=int(b["first"]))
    same(g, b, "first_entity offset")


def test_append_to_scene_offsets_ids():
    from engine_b200 import synth

    s = synth.uniform_boxes(5)
    data = make_glb(3, n_nodes=15, n_meshes=4)
    root = gltf.append_to_scene(s, data, stem="file")
    g = gltf.load_colliders(data, first_entity=root, stem="file")
    assert root == 6 and s.n_entities == 5 + len(g["ids"]) and s.parent[root] == 0
    assert (s.parent[root + 1:] >= root).all() and (s.parent[root + 1:] < np.arange(root + 1, s.n_entities + 1)).all(), "parents precede children"
    assert (s.collider_entity[5:] == g["collider_entity"]).all()


def test_malformed_inputs_fail_cleanly():
    good = make_glb(1)
    for bad, status in [
        (b"", gltf.ERR_FORMAT),
        (b"glTF" + bytes(16), gltf.ERR_FORMAT),
        (b"gltf" + good[4:], gltf.ERR_FORMAT),
        (good[:40], gltf.ERR_FORMAT),                      # truncated JSON chunk
        (good[:20] + b"{" * 16 + good[36:], gltf.ERR_FORMAT),  # broken JSON
    ]:
        with pytest.raises(gltf.GltfError) as e:
            gltf.load_colliders(bad)
        assert e.value.status == status
    with pytest.raises(gltf.GltfError) as e:
        gltf.load_colliders("/nonexistent/file.glb")
    assert e.value.status == gltf.ERR_IO
    # a node cycle must not hang
    import json
    import struct

    doc = {"asset": {"version": "2.0"}, "scenes": [{"nodes": [0]}], "nodes": [{"children": [1]}, {"children": [0]}]}
    js = json.dumps(doc).encode()
    js += b" " * (-len(js) % 4)
    glb = b"glTF" + struct.pack("<II", 2, 20 + len(js)) + struct.pack("<II", len(js), 0x4E4F534A) + js
    with pytest.raises(gltf.GltfError):
        gltf.load_colliders(glb)
