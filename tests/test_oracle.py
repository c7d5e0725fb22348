"""CPU: the oracle port (oracle/port.c) against the golden vectors produced by the real reference,
and — where oracle/_ref/libref_engine.so exists — against the reference itself on fresh inputs."""
import numpy as np
import pytest

from conftest import SCENE_FIXTURES, assert_raycast_equal, load_golden, GOLDEN
from engine_b200 import synth
from oracle.oracle import NODE_DTYPE, RAYCAST_DTYPE, RefEngine, TYPE_BV, TYPE_EMPTY, TYPE_ENTITY
import os


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_pod_layouts(port):
    assert RAYCAST_DTYPE.itemsize == 36 and NODE_DTYPE.itemsize == 60
    if RefEngine.available():
        r = RefEngine()
        assert [r.lib.refh_sizeof(i) for i in range(6)] == [24, 36, 24, 60, 144, 40]


def test_kat_ray_box(port):
    z = np.load(os.path.join(GOLDEN, "kat.npz"))
    for i in range(len(z["rays"])):
        hit, t, side = port.ray_box(z["rays"][i], z["boxes"][i])
        assert hit == bool(z["hit"][i]), i
        # tmin is compared numerically: NaN == NaN positions, +-0 equal (see conftest.assert_raycast_equal)
        assert (t == z["t"][i]) or (np.isnan(t) and np.isnan(z["t"][i])), (i, t, z["t"][i])
        assert side == z["side"][i], i


def test_kat_transform_box_and_normalize(port):
    z = np.load(os.path.join(GOLDEN, "kat.npz"))
    for i in range(len(z["tb_in"])):
        out = port.transform_box(z["tb_in"][i], z["tb_mat"][i])
        assert (bits(out) == bits(z["tb_out"][i])).all(), i
    for i in range(len(z["norm_in"])):
        assert (bits(port.normalize(z["norm_in"][i])) == bits(z["norm_out"][i])).all(), i


@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_golden_scene(port, name):
    g = load_golden(name)
    s = g["scene"]
    world = port.transform_update(s.parent, s.pos, s.rot, s.scl)
    assert (bits(world[1:]) == bits(g["world"])).all(), "world matrices must be bit-exact"
    boxes = port.refit(s.collider_entity, s.collider_aabb, world)
    assert (g["world_aabb_entity"] == s.collider_entity).all()
    assert (bits(boxes.view(np.float32).reshape(-1, 6)) == bits(g["world_aabb"])).all(), "world AABBs must be bit-exact"
    # reference traversal restated, on the reference's own tree
    out = port.raycast_tree(g["bvh"], g["rays"])
    assert_raycast_equal(out, g["raycast"], name + " tree")
    # brute force over leaves in the tree's DFS order == the tree (incl. tie winners)
    ent, box = port.tree_leaves_dfs(g["bvh"])
    assert sorted(ent.tolist()) == sorted(s.collider_entity.tolist())
    out_b, ties = port.raycast_brute(ent, box, g["rays"], nthreads=4, want_ties=True)
    assert_raycast_equal(out_b, g["raycast"], name + " brute")
    # a foreign topology only changes hit_entity (and with it the normal) on exact-tie rays
    foreign = port.build_binned_sah(s.collider_entity, boxes)
    out_f = port.raycast_tree(foreign, g["rays"])
    assert_raycast_equal(out_f, g["raycast"], name + " foreign tree", ignore_entity_on=ties > 0)


def test_reference_node_shapes(port):
    """Trees from BuildNode only contain {BV,BV}, {Entity,Entity}, {Entity,Empty} nodes, root last
    (reference src/system_collisions.cpp:351-369,372); the binned builder must emit the same shapes."""
    g = load_golden("game600")
    boxes = port.refit(g["scene"].collider_entity, g["scene"].collider_aabb, port.transform_update(*[getattr(g["scene"], k) for k in ("parent", "pos", "rot", "scl")]))
    for tree in (g["bvh"], port.build_binned_sah(g["scene"].collider_entity, boxes)):
        pairs = set(zip(tree["type1"].tolist(), tree["type2"].tolist()))
        assert pairs <= {(TYPE_BV, TYPE_BV), (TYPE_ENTITY, TYPE_ENTITY), (TYPE_ENTITY, TYPE_EMPTY)}
        bv = tree["type1"] == TYPE_BV
        assert (tree["index1"][bv] < np.nonzero(bv)[0]).all() and (tree["index2"][bv] < np.nonzero(bv)[0]).all()  # post-order


def test_empty_scene_misses(port):
    out = port.raycast_tree(np.zeros(0, NODE_DTYPE), synth.random_rays(8, extent=5))
    assert not out.view(np.uint8).any()  # value-initialised Raycast (reference :164-167)


@pytest.mark.skipif(not RefEngine.available(), reason="reference sources not present (GPU box): golden fixtures cover this")
@pytest.mark.parametrize("seed", [21, 22])
def test_port_vs_live_reference(port, seed):
    s = synth.hierarchy_scene(800, 12, seed=seed, extent=40, negative_scale_fraction=0.2)
    rays = np.concatenate([synth.random_rays(3000, seed=seed, extent=50, normalised=False), synth.axis_rays(1000, seed=seed, extent=40)])
    r = RefEngine()
    r.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
    r.add_colliders(s.collider_entity, s.collider_aabb)
    r.scene_update()
    world = port.transform_update(s.parent, s.pos, s.rot, s.scl)
    assert (bits(world[1:]) == bits(r.world_matrices(np.arange(1, s.n_entities + 1, dtype=np.uint32)))).all()
    ent, box, _ = r.refit(s.n_colliders)
    assert (bits(port.refit(s.collider_entity, s.collider_aabb, world).view(np.float32)) == bits(box.view(np.float32))).all()
    want = r.raycast(rays)
    assert_raycast_equal(port.raycast_tree(r.bvh(), rays), want, "port vs live reference")
    # inject a foreign tree into the reference's public bvh_ and let ITS GetRaycast walk it
    foreign = port.build_binned_sah(ent, box)
    r.set_bvh(foreign)
    inj = r.raycast(rays, nthreads=2)
    assert_raycast_equal(port.raycast_tree(foreign, rays), inj, "port vs reference on injected tree")
    e2, b2 = port.tree_leaves_dfs(foreign)
    assert_raycast_equal(port.raycast_brute(e2, b2, rays, nthreads=4), inj, "brute(DFS order) vs reference on injected tree")
