"""CPU: the LOGIC of the GPU kernels (per-thread bodies from engine_b200/csrc/*.cuh executed on the
host by tests/emu) against the oracle and the reference's golden vectors.  The CUDA library itself is
tested by the -m gpu tests; this file only guards the algorithms while no GPU is attached."""
import numpy as np
import pytest

from conftest import SCENE_FIXTURES, assert_raycast_equal, load_golden
from emu_loader import LEAF_BIT, Emu
from engine_b200 import synth
from oracle.oracle import TYPE_BV, TYPE_EMPTY, TYPE_ENTITY, RefEngine


@pytest.fixture(scope="module")
def emu():
    return Emu()


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def check_tree(tree):
    """Structural invariants of the LBVH: every leaf once, DFS leaf order == Morton order, parent
    boxes are the exact union of their children."""
    n, nodes, root = tree["n"], tree["nodes"], tree["root"]
    if n == 1:
        assert nodes["ref0"][0] == LEAF_BIT and nodes["ref1"][0] == 0xFFFFFFFF and np.isnan(nodes["lo1"][0]).all()
        return 1
    wa = tree["world_aabb"][tree["sorted_prim"]]
    order = []
    depth_max = 0

    def rec(p, d):
        nonlocal depth_max
        depth_max = max(depth_max, d)
        lo = np.full(3, np.inf, np.float32)
        hi = np.full(3, -np.inf, np.float32)
        first = len(order)
        for s in (0, 1):
            ref = int(nodes[f"ref{s}"][p])
            if ref & LEAF_BIT:
                i = ref & ~LEAF_BIT
                order.append(i)
                assert (bits(nodes[f"lo{s}"][p]) == bits(wa[i, :3])).all() and (bits(nodes[f"hi{s}"][p]) == bits(wa[i, 3:])).all()
                clo, chi = wa[i, :3], wa[i, 3:]
            else:
                clo, chi = rec(ref, d + 1)
                assert (nodes[f"lo{s}"][p] == clo).all() and (nodes[f"hi{s}"][p] == chi).all()
            lo, hi = np.minimum(lo, clo), np.maximum(hi, chi)
        # aux0 / aux1 = first / last leaf position of the node's range
        assert nodes["aux0"][p] == order[first] and nodes["aux1"][p] == order[-1]
        return lo, hi

    import sys

    sys.setrecursionlimit(10000)
    rec(root, 1)
    assert order == list(range(n)), "depth-first leaf order must equal Morton order"
    assert (np.diff(tree["sorted_keys"].astype(np.uint64)) >= 0).all() or (tree["sorted_keys"][1:] >= tree["sorted_keys"][:-1]).all()
    return depth_max


@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_emu_against_golden(emu, port, name):
    g = load_golden(name)
    s = g["scene"]
    world = emu.propagate(s.parent, s.pos, s.rot, s.scl)
    assert (bits(world[1:]) == bits(g["world"])).all(), "world matrices bit-exact vs reference"
    boxes = emu.refit(s.collider_entity, s.collider_aabb, world)
    assert (bits(boxes) == bits(g["world_aabb"])).all(), "world AABBs bit-exact vs reference"
    # reference-topology traversal on the reference's own tree: everything incl. tie winners
    assert_raycast_equal(emu.trace_reftopo(g["bvh"], g["rays"]), g["raycast"], name + " reftopo")
    # production tree
    tree = emu.build(boxes, s.collider_entity)
    check_tree(tree)
    got = emu.trace_lbvh(tree, g["rays"])
    # (1) == brute force with the tie rule "last in Morton order wins"
    brute, ties = port.raycast_brute(tree["leaf_entity"], boxes[tree["sorted_prim"]], g["rays"], nthreads=4, want_ties=True)
    assert_raycast_equal(got, brute, name + " lbvh vs brute")
    # (2) == the reference's recursive traversal walking OUR exported tree
    exported = emu.export(tree)
    assert exported["type1"][-1] != TYPE_EMPTY
    pairs = set(zip(exported["type1"].tolist(), exported["type2"].tolist()))
    assert pairs <= {(TYPE_BV, TYPE_BV), (TYPE_ENTITY, TYPE_ENTITY), (TYPE_ENTITY, TYPE_EMPTY)}
    assert_raycast_equal(got, port.raycast_tree(exported, g["rays"]), name + " lbvh vs reference traversal of exported tree")
    # (3) == the reference on ITS OWN tree except hit_entity/normal on exact ties
    assert_raycast_equal(got, g["raycast"], name + " lbvh vs reference", ignore_entity_on=ties > 0)
    # reftopo kernel body on the exported tree gives the same again
    assert_raycast_equal(emu.trace_reftopo(exported, g["rays"]), got, name + " reftopo on exported")


def test_emu_thread_order_independence(emu):
    s = synth.uniform_boxes(3000, seed=31, extent=40)
    world = emu.propagate(s.parent, s.pos, s.rot, s.scl)
    boxes = emu.refit(s.collider_entity, s.collider_aabb, world)
    a = emu.build(boxes, s.collider_entity)
    rng = np.random.default_rng(5)
    for _ in range(3):
        b = emu.build(boxes, s.collider_entity, order=rng.permutation(3000).astype(np.uint32))
        assert a["root"] == b["root"] and (a["nodes"].view(np.uint8) == b["nodes"].view(np.uint8)).all()


@pytest.mark.parametrize("bits_per_axis", [4, 10, 16, 21])
def test_emu_duplicate_keys_and_ties(emu, port, bits_per_axis):
    """Coincident and grid-aligned boxes: duplicate Morton keys (position tie-break in the hierarchy)
    and exact tmin ties (last-in-Morton-order rule)."""
    s = synth.grid_boxes(6, 6, 6, spacing=3.0)
    world = emu.propagate(s.parent, s.pos, s.rot, s.scl)
    boxes = emu.refit(s.collider_entity, s.collider_aabb, world)
    boxes = np.concatenate([boxes, boxes[:50], boxes[:50]])  # exact duplicates
    ent = np.arange(1, boxes.shape[0] + 1, dtype=np.uint32)
    tree = emu.build(boxes, ent, bits=bits_per_axis)
    depth = check_tree(tree)
    assert depth <= 96
    rays = synth.axis_rays(5000, extent=18)
    got = emu.trace_lbvh(tree, rays)
    brute, ties = port.raycast_brute(tree["leaf_entity"], boxes[tree["sorted_prim"]], rays, nthreads=4, want_ties=True)
    assert (ties > 0).sum() > 50
    assert_raycast_equal(got, brute, "ties")
    assert_raycast_equal(got, port.raycast_tree(emu.export(tree), rays), "ties via exported tree")


def test_emu_all_identical_boxes(emu, port):
    boxes = np.tile(np.array([[0, 0, 0, 1, 1, 1]], np.float32), (257, 1))
    ent = np.arange(1, 258, dtype=np.uint32)
    tree = emu.build(boxes, ent)
    check_tree(tree)
    rays = synth.random_rays(500, extent=3, normalised=False)
    got = emu.trace_lbvh(tree, rays)
    assert (got["hit_entity"][got["hit"] > 0] == 257).all()  # last in order wins every tie
    assert_raycast_equal(got, port.raycast_tree(emu.export(tree), rays), "identical boxes")


@pytest.mark.skipif(not RefEngine.available(), reason="reference sources not present")
def test_emu_export_walked_by_live_reference(emu):
    s = synth.hierarchy_scene(2500, 30, seed=41, extent=60)
    rays = np.concatenate([synth.random_rays(4000, seed=42, extent=70, normalised=False), synth.axis_rays(500, seed=43, extent=60)])
    world = emu.propagate(s.parent, s.pos, s.rot, s.scl)
    boxes = emu.refit(s.collider_entity, s.collider_aabb, world)
    tree = emu.build(boxes, s.collider_entity)
    r = RefEngine()
    r.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
    r.add_colliders(s.collider_entity, s.collider_aabb)
    r.transform_update()
    r.set_bvh(emu.export(tree))  # the reference's public bvh_ member
    assert_raycast_equal(emu.trace_lbvh(tree, rays), r.raycast(rays), "reference GetRaycast on exported tree")
