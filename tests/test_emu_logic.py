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


# ---------------------------------------------------------------------------------------------
# Morton bit allocation (make_morton_plan / morton_key): the K key bits follow the scene's extents
# ---------------------------------------------------------------------------------------------
def _classic_interleave(q, bits):
    key = np.zeros(q.shape[0], np.uint64)
    for b in range(bits):
        for k in range(3):
            key |= ((q[:, k].astype(np.uint64) >> np.uint64(b)) & np.uint64(1)) << np.uint64(3 * b + (2 - k))
    return key


def test_morton_plan_cubic_scene_is_the_plain_interleave(emu):
    rng = np.random.default_rng(1)
    c = rng.uniform(-100, 100, (5000, 3)).astype(np.float32)
    cmin, cmax = c.min(0), c.max(0)
    for key_bits, bits in ((32, 11), (30, 10), (40, 14), (48, 16), (63, 21), (24, 8), (31, 11), (26, 9), (25, 9), (33, 11), (16, 6)):  # <= 32 bits: the 32-bit path, every `drop`
        plan, keys = emu.morton(c, cmin, cmax, key_bits)
        assert plan["d2"] == 0 and plan["shift_a"] == 0 and plan["nc"] == bits and plan["drop"] == 3 * bits - key_bits
        order = [plan["a"], plan["b"], plan["c"]]
        sinv = np.float32(1 << bits) / (cmax - cmin)
        q = np.clip((c - cmin) * sinv, 0, np.float32((1 << bits) - 1)).astype(np.uint32)
        want = _classic_interleave(q[:, order], bits) >> np.uint64(plan["drop"])
        assert (keys == want).all()
        assert int(keys.max()) < (1 << key_bits)


@pytest.mark.parametrize("ext,key_bits", [((1000, 1000, 50), 32), ((50, 1000, 1000), 32), ((1000, 3, 1000), 40), ((4000, 1, 1), 32), ((1, 1, 0), 32),
                                          ((7, 0, 0), 32), ((1e6, 1e-3, 1.0), 48), ((1000, 900, 50), 63), ((3, 2, 1), 12), ((1000, 1000, 50), 48)])
def test_morton_plan_follows_the_extents(emu, ext, key_bits):
    rng = np.random.default_rng(2)
    ext = np.asarray(ext, np.float32)
    c = (rng.uniform(0, 1, (20000, 3)) * ext).astype(np.float32)
    c[0], c[1] = 0, ext  # the bounds are attained
    cmin, cmax = c.min(0), c.max(0)
    plan, keys = emu.morton(c, cmin, cmax, key_bits)
    axes = [plan["a"], plan["b"], plan["c"]]
    assert sorted(axes) == [0, 1, 2]
    e = (cmax - cmin)[axes]
    assert e[0] >= e[1] >= e[2]
    nc, d2 = plan["nc"], plan["d2"]
    total = 3 * nc + 2 * d2
    assert int(keys.max()) < (1 << key_bits)
    assert plan["drop"] in (0, 1, 2)
    # bits per axis: a gets the rest of the key
    na = key_bits + plan["drop"] - total + nc + d2
    nb, n_c = nc + d2, nc
    assert na >= nb >= n_c >= 0 and na + nb + n_c == key_bits + plan["drop"]
    # cells as cubic as powers of two allow: wherever two axes both received bits and both have an extent,
    # their cell sizes are within a factor 2 * sqrt(2) of each other (sqrt 2 from the rounding, 2 from the dropped bits)
    cells = [e[0] / 2.0**na, e[1] / 2.0**nb, e[2] / 2.0**n_c]
    got_bits = [na, nb, n_c]
    for i in range(3):
        for j in range(i + 1, 3):
            if e[i] > 0 and e[j] > 0 and got_bits[j] > 0 and got_bits[i] < 21 and got_bits[j] < 21:
                r = cells[i] / cells[j]
                assert 1 / 2.9 < r < 2.9, (plan, cells)
    # an axis that received no bit must be the thinnest: its whole extent is no larger than ~1.5 cells of the others
    if e[0] > 0 and na <= 21:
        for j in (1, 2):
            if got_bits[j] == 0:
                assert e[j] <= cells[0] * 1.5
    # the key is monotone along every axis (other coordinates fixed)
    base = c[2:12]
    for k in range(3):
        if ext[k] == 0:
            continue
        line = np.repeat(base, 50, axis=0)
        line[:, k] = np.tile(np.linspace(0, ext[k], 50, dtype=np.float32), base.shape[0])
        _, lk = emu.morton(line, cmin, cmax, key_bits)
        lk = lk.reshape(-1, 50).astype(np.int64) if key_bits < 63 else lk.reshape(-1, 50)
        assert (lk[:, 1:] >= lk[:, :-1]).all()
    # sorting by the key clusters space: with extents in all three axes, neighbours in key order are close
    if (ext > 0).all() and key_bits >= 30:
        o = np.argsort(keys, kind="stable")
        d = np.linalg.norm(np.diff(c[o].astype(np.float64), axis=0), axis=1)
        rand = np.linalg.norm(np.diff(c.astype(np.float64), axis=0), axis=1)
        assert np.median(d) < 0.2 * np.median(rand)


def test_morton_plan_flat_world_example(emu):
    """BASELINE config 2's world: 1000 x 1000 x 50 -> z gets 4 bits fewer than x and y (50 * 2^4 = 800 ~ 1000)."""
    c = np.array([[0, 0, 0], [1000, 1000, 50]], np.float32)
    plan, _ = emu.morton(c, c.min(0), c.max(0), 32)
    assert plan["c"] == 2 and plan["d2"] == 4 and plan["nc"] == 8 and plan["drop"] == 0 and plan["shift_a"] == 0  # x, y: 12 bits, z: 8


@pytest.mark.parametrize("with_nan", [False, True])
def test_emu_degenerate_colliders_and_rays(emu, port, with_nan):
    """NaN / inf / huge / zero-scale / inverted colliders; the kernel bodies must return what the reference's own
    arithmetic returns leaf by leaf (brute force with its slab test) — wild leaves are always entered (b2col_bvh.cuh WILD_BIT) —
    for sane and for degenerate rays, including the rays whose NaN origin coordinate makes the reference ignore an axis."""
    s = synth.degenerate_scene(5000, with_nan=with_nan)
    world = emu.propagate(s.parent, s.pos, s.rot, s.scl)
    assert (bits(world) == bits(port.transform_update(s.parent, s.pos, s.rot, s.scl))).all()
    boxes = emu.refit(s.collider_entity, s.collider_aabb, world)
    assert (bits(boxes) == bits(port.refit(s.collider_entity, s.collider_aabb, world).view(np.float32).reshape(-1, 6))).all()
    assert (~np.isfinite(boxes)).any(axis=1).sum() >= 10
    tree = emu.build(boxes, s.collider_entity)
    assert tree["has_wild"] == 1
    rays = np.concatenate([synth.degenerate_rays(4000), synth.random_rays(4000, seed=5, extent=70)])
    got = emu.trace_lbvh(tree, rays)
    brute = port.raycast_brute(tree["leaf_entity"], boxes[tree["sorted_prim"]], rays, nthreads=8)
    assert_raycast_equal(got, brute, "degenerate scene vs brute force", nan_equal=True)
    # without the boxes that swallow every ray (huge on all three axes: infinite / 1e30 scale) the other kinds decide
    with np.errstate(invalid="ignore"):
        keep = ~((boxes[:, 3:] - boxes[:, :3]) > 1e6).all(axis=1)
    tree2 = emu.build(boxes[keep], s.collider_entity[keep])
    got2 = emu.trace_lbvh(tree2, rays)
    brute2 = port.raycast_brute(tree2["leaf_entity"], boxes[keep][tree2["sorted_prim"]], rays, nthreads=8)
    assert_raycast_equal(got2, brute2, "degenerate scene without the all-swallowing boxes vs brute force", nan_equal=True)
    if not with_nan:
        assert 100 < got2["hit"].sum() < rays.shape[0] - 100  # no NaN plane: unions are exact, so the reference's own walk of the exported tree agrees too
        assert_raycast_equal(got2, port.raycast_tree(emu.export(tree2), rays, nthreads=8), "exported", nan_equal=True)


def test_shard_bounds_match_python():
    from engine_b200 import collision, dist

    for n in (0, 1, 7, 1000, 1001, 16 * 1024 * 1024, (1 << 40) + 12345):
        for world in (1, 2, 3, 8):
            cover = 0
            for r in range(world):
                lo, hi = collision.shard_bounds(n, world, r)
                assert (lo, hi) == dist.shard_bounds(n, world, r)
                assert lo == cover
                cover = hi
            assert cover == n
