"""Generate tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/libref_engine.so,
compiled from /root/reference by oracle/Makefile).  Run in the build container only:

    python tests/golden/make_golden.py

Each fixture holds the inputs (scene arrays, rays) and the reference's own outputs: world
matrices (TransformSystem::onUpdate), world AABBs (transformBox via the refit loop), the reference
BVH (CollisionSystem::bvh_ after Scene::Update) and Raycast results (CollisionSystem::GetRaycast).
The reference ships no tests or golden vectors of its own, so these are the pins.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from engine_b200 import synth  # noqa: E402
from oracle.oracle import RefEngine  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def load_ref(s):
    r = RefEngine()
    ids = r.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
    assert ids[0] == 1
    r.add_colliders(s.collider_entity, s.collider_aabb)
    return r


def scene_fixture(name, s, rays):
    r = load_ref(s)
    r.scene_update()
    world = r.world_matrices(np.arange(1, s.n_entities + 1, dtype=np.uint32))
    ent, box, _ = r.refit(s.n_colliders)
    bvh = r.bvh()
    out = r.raycast(rays)
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        parent=s.parent, pos=s.pos, rot=s.rot, scl=s.scl, collider_entity=s.collider_entity, collider_aabb=s.collider_aabb,
        world=world, world_aabb_entity=ent, world_aabb=box.view(np.float32).reshape(-1, 6),
        bvh=bvh.view(np.uint8).reshape(-1, 60), rays=rays, raycast=out.view(np.uint8).reshape(-1, 36),
    )
    print(name, "colliders", s.n_colliders, "nodes", len(bvh), "rays", len(rays), "hits", int(out["hit"].sum()))


def kat_fixture():
    """Known-answer vectors for the file-static helpers RayBoxIntersection (:62-100), transformBox (:14-41)
    and glm::normalize as GetRaycast uses it (:162)."""
    r = RefEngine()
    rng = np.random.default_rng(11)
    inf = np.float32(np.inf)
    rays, boxes = [], []
    unit = [-1, -1, -1, 1, 1, 1]
    hand = [
        ([0, 0, 5, 0, 0, -1], unit),        # the game's fall ray: +-inf reciprocals on x,y
        ([0, 0, 5, 0.0, -0.0, -1], unit),   # -0 component -> -inf reciprocal
        ([1, 0, 5, 0, 0, -1], unit),        # origin on the max.x plane with d.x = 0 -> NaN dropped by fminf/fmaxf
        ([-1, 0, 5, 0, 0, -1], unit),       # origin on the min.x plane
        ([0, 0, 0, 1, 0, 0], unit),         # origin inside: negative tmin is a hit
        ([0, 0, 0, 0.6, 0.8, 0], unit),
        ([-5, 0, 0, 1, 0, 0], unit),        # Left face
        ([5, 0, 0, -1, 0, 0], unit),        # Right face
        ([0, -5, 0, 0, 1, 0], unit),        # Front
        ([0, 5, 0, 0, -1, 0], unit),        # Back
        ([0, 0, -5, 0, 0, 1], unit),        # Bottom
        ([-5, -5, -5, 1, 1, 1], unit),      # exact corner: side chain picks Left first
        ([5, 5, 5, -1, -1, -1], unit),      # opposite corner: Right
        ([-5, -5, 0, 1, 1, 0], unit),       # edge
        ([-2000, 0, 0, 1, 0, 0], unit),     # tmin >= 1000 -> miss
        ([-1001, 0, 0, 1, 0, 0], unit),     # tmin == 1000 exactly -> miss (strict <)
        ([-1000.5, 0, 0, 1, 0, 0], unit),
        ([5, 0, 0, 1, 0, 0], unit),         # box behind the ray
        ([0, 3, 0, 1, 0, 0], unit),         # parallel, outside the slab
        ([0, 0, 0, 1, 0, 0], [0, 0, 0, 0, 0, 0]),   # degenerate box at the origin
        ([0, 0, 0, 0, 0, 1], [0, 0, 1, 0, 0, 1]),
        ([0.5, 0.5, 3, 0, 0, -1], [0, 0, 0, 1, 1, 0.01]),  # thin land tile
    ]
    for ry, bx in hand:
        rays.append(ry)
        boxes.append(bx)
    for _ in range(2000):
        c = rng.uniform(-10, 10, 3)
        h = rng.uniform(0.01, 4, 3)
        o = rng.uniform(-15, 15, 3)
        kind = rng.integers(0, 4)
        if kind == 0:
            d = rng.standard_normal(3)
            d /= np.linalg.norm(d)
        elif kind == 1:
            d = np.zeros(3)
            d[rng.integers(0, 3)] = rng.choice([-1.0, 1.0])
            o = np.round(o)
            c, h = np.round(c), np.maximum(np.round(h), 1)
        elif kind == 2:
            d = np.array([np.cos(1.0 * _), np.sin(1.0 * _), 0.0])
        else:
            t = (c - o)
            d = t / np.linalg.norm(t)
        rays.append(list(o) + list(d))
        boxes.append(list(c - h) + list(c + h))
    rays = np.array(rays, np.float32)
    boxes = np.array(boxes, np.float32)
    hit = np.zeros(len(rays), np.uint8)
    t = np.zeros(len(rays), np.float32)
    side = np.zeros(len(rays), np.int32)
    for i in range(len(rays)):
        hit[i], t[i], side[i] = r.ray_box(rays[i], boxes[i])
    # transformBox: random TRS-like and arbitrary matrices
    mats = rng.standard_normal((500, 16)).astype(np.float32) * 3
    mats[:, 3] = 0; mats[:, 7] = 0; mats[:, 11] = 0; mats[:, 15] = 1
    mats[0] = np.eye(4, dtype=np.float32).reshape(-1)
    tb_in = np.sort(rng.uniform(-5, 5, (500, 2, 3)).astype(np.float32), axis=1).reshape(500, 6)
    tb_out = np.stack([r.transform_box(tb_in[i], mats[i]) for i in range(500)])
    nv = (rng.standard_normal((500, 3)) * rng.choice([1e-3, 1.0, 1e3], (500, 1))).astype(np.float32)
    nv_out = np.stack([r.normalize(v) for v in nv])
    np.savez_compressed(os.path.join(OUT, "kat.npz"), rays=rays, boxes=boxes, hit=hit, t=t, side=side,
                        tb_mat=mats, tb_in=tb_in, tb_out=tb_out, norm_in=nv, norm_out=nv_out)
    print("kat", len(rays), "ray/box pairs,", int(hit.sum()), "hits;", len(mats), "transformBox;", len(nv), "normalize")


def gltf_fixture():
    """Collider ingestion: what the UNMODIFIED reference loadGLTF (src/gltf_loader.cpp) creates
    for (a) synthetic .glb files from tests/gltf_synth.py — the bytes are stored too — and (b) the reference's own
    models under test/res/models (only the derived entity / collider arrays are stored, not the assets)."""
    import glob
    import tempfile

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from gltf_synth import make_glb

    out = {}
    names = []

    def record(key, path):
        try:
            g = RefEngine().load_gltf(path)
        except RuntimeError as ex:
            out[key + "/error"] = np.frombuffer(str(ex).encode(), np.uint8)
            print(key, "reference throws:", ex)
            return
        for k in ("parent", "pos", "rot", "scl", "collider_entity", "collider_aabb"):
            out[f"{key}/{k}"] = g[k]
        out[key + "/tags"] = np.frombuffer("\n".join(g["tags"]).encode(), np.uint8)
        print(key, "entities", len(g["ids"]), "colliders", len(g["collider_entity"]))

    for seed in range(12):
        data = make_glb(seed, n_nodes=6 + 3 * seed, n_meshes=1 + seed % 5, flavour=["mixed", "single_scene"][seed % 2])
        key = f"synth{seed}"
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, key + ".glb")
            open(path, "wb").write(data)
            record(key, path)
        out[key + "/glb"] = np.frombuffer(data, np.uint8)
        names.append(key)
    for path in sorted(glob.glob("/root/reference/test/res/models/*.glb")):
        key = "model_" + os.path.basename(path)[:-4]
        record(key, path)
        names.append(key)
    out["names"] = np.frombuffer("\n".join(names).encode(), np.uint8)
    np.savez_compressed(os.path.join(OUT, "gltf.npz"), **out)


def game_fixture():
    """The test game's player controller: 600 frames of the UNMODIFIED reference
    test/src/camera_controller.cpp + reference CollisionSystem under scripted input."""
    s = synth.game_scene(140)
    r = load_ref(s)
    r.add_camera((8.0, 2.0, 4.0))
    inp = synth.game_inputs(600)
    state = r.run_game(inp)
    np.savez_compressed(os.path.join(OUT, "game_traj.npz"), input=inp, state=state, camera=np.array([8.0, 2.0, 4.0], np.float32))
    print("game_traj frames", len(inp), "grounded", int(state[:, 12].sum()))


if __name__ == "__main__":
    if "--only-new" in sys.argv:
        gltf_fixture()
        game_fixture()
        sys.exit(0)
    gltf_fixture()
    game_fixture()
    kat_fixture()
    scene_fixture("game600", synth.game_scene(600), np.concatenate([synth.game_rays(100), synth.random_rays(2000, extent=100, normalised=False)]))
    scene_fixture("hier1500", synth.hierarchy_scene(1500, 20, extent=50, negative_scale_fraction=0.1), synth.random_rays(4000, extent=60, normalised=False))
    scene_fixture("grid512", synth.grid_boxes(8, 8, 8), synth.axis_rays(6000, extent=16))
    scene_fixture("city2000", synth.static_city(2000), synth.camera_rays(80, 60, eye=(0, 0, 60)))
    scene_fixture("single", synth.uniform_boxes(1, extent=1.0), synth.random_rays(64, extent=5, normalised=False))
    scene_fixture("pair", synth.uniform_boxes(2, extent=2.0), synth.random_rays(64, extent=5, normalised=False))
    scene_fixture("triple", synth.uniform_boxes(3, extent=3.0), synth.random_rays(64, extent=5, normalised=False))
