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


def dequant(tree):
    """World-space boxes of both slots of every quantised node (15-bit cells, flag bit 0x8000)."""
    g = tree["grid"][0]
    w = tree["qnodes"]["w"].astype(np.uint64)
    assert ((w[:, :6] & 0x8000) != 0).all() and (((w[:, :6] >> 16) & 0x8000) != 0).all()
    lo = (w[:, :6] & 0x7FFF).astype(np.float64)          # (n, 6): slot0 xyz, slot1 xyz
    hi = ((w[:, :6] >> 16) & 0x7FFF).astype(np.float64)
    gm, cell = g["gmin"].astype(np.float64), g["cell"].astype(np.float64)
    return {s: np.concatenate([gm + lo[:, 3 * s : 3 * s + 3] * cell, gm + hi[:, 3 * s : 3 * s + 3] * cell], 1) for s in (0, 1)}


@pytest.mark.parametrize("extent,offset,box", [(1000.0, 0.0, 2.0), (1.0, 0.0, 0.01), (1e-3, 0.0, 1e-5), (50.0, 1.0e6, 1.0), (1.0e5, -3.0e5, 30.0), (20.0, 0.0, 15.0)])
def test_emu_quantised_nodes_are_conservative(emu, port, extent, offset, box):
    """The 16-bit grid boxes must contain the exact boxes (they only prune), for scenes of any scale and
    position, and the quantised traversal must still equal brute force — including rays that start far
    outside the grid (provable-miss shortcut) and rays lying exactly on box planes."""
    rng = np.random.default_rng(int(extent * 7) % 1000 + 3)
    n = 3000
    c = (rng.uniform(-extent, extent, (n, 3)) + offset).astype(np.float32)
    h = rng.uniform(0.1 * box, box, (n, 3)).astype(np.float32)
    boxes = np.concatenate([c - h, c + h], 1).astype(np.float32)
    ent = np.arange(1, n + 1, dtype=np.uint32)
    tree = emu.build(boxes, ent)
    dq = dequant(tree)
    nodes = tree["nodes"]
    for s in (0, 1):
        lo, hi = nodes[f"lo{s}"].astype(np.float64), nodes[f"hi{s}"].astype(np.float64)
        assert (dq[s][:, :3] <= lo).all() and (dq[s][:, 3:] >= hi).all(), "grid box does not contain the exact box"
    g = tree["grid"][0]
    assert (g["cell"] >= 1000.0 * 2.0**-18 * 0.999).all()
    o = (rng.uniform(-1.2 * extent, 1.2 * extent, (6000, 3)) + offset).astype(np.float32)
    d = rng.standard_normal((6000, 3)).astype(np.float32)
    d[:1500] = np.sign(d[:1500]) * (np.abs(d[:1500]) > 1.0)          # axis-parallel / diagonal / zero
    o[1500:2500] = boxes[rng.integers(0, n, 1000), :3]               # origins exactly on box min planes
    o[2500:3000] += np.float32(40000.0) * np.sign(d[2500:3000])      # far outside: out of the grid's range
    o[3000:3200] = np.nan
    rays = np.concatenate([o, d], 1).astype(np.float32)
    got = emu.trace_lbvh(tree, rays)
    brute = port.raycast_brute(tree["leaf_entity"], boxes[tree["sorted_prim"]], rays, nthreads=4)
    assert_raycast_equal(got, brute, f"quantised traversal, extent {extent} offset {offset}")
    assert got["hit"].sum() > 0
