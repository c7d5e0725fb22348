"""GPU (-m gpu): the CUDA path, called through the C ABI (engine_b200.collision -> libb2col.so),
against the oracle: golden vectors from the real reference, the oracle port, the live reference where
oracle/_ref/libref_engine.so travelled with the snapshot, and size-independent properties at
BASELINE.json's full sizes.  Parity bar: hit flag + hit_entity bit-exact; distance / location /
normal held to exact float equality (stricter than the 1e-5 relative tolerance of north_star)."""
import os

import numpy as np
import pytest
import torch

from conftest import SCENE_FIXTURES, assert_raycast_equal, load_golden
from emu_loader import Emu
from engine_b200 import capi, synth
from oracle.oracle import Dropin2Engine, DropinEngine, RefEngine

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def CollisionSystem():
    if not torch.cuda.is_available():
        pytest.fail("no CUDA device: the gpu-marked tests must run on the B200 box")
    from engine_b200.collision import CollisionSystem as CS

    return CS


@pytest.fixture(scope="module")
def emu():
    return Emu()


_REF_WALKER = None


def reference_walk(exported, rays, port):
    """GetRaycast of the LIVE reference (oracle/_ref/libref_engine.so, unmodified src/system_collisions.cpp:160-199,379-540) over
    `rays`, walking `exported` assigned to its public bvh_ member, on all host threads — every ray, no sampling.  The pinned C port
    (second opinion, and the checker where the reference library did not travel) takes over when the library is absent."""
    global _REF_WALKER
    threads = os.cpu_count() or 8
    if RefEngine.available():
        if _REF_WALKER is None:
            _REF_WALKER = RefEngine()
        _REF_WALKER.set_bvh(exported)
        return _REF_WALKER.raycast(rays, nthreads=threads)
    return port.raycast_tree(exported, rays, nthreads=threads)


def read_blob(cs, n):
    """(qnodes [n-1, 8] u32, leaf records [n, 8] u32) of the tree currently in the context's BVH blob"""
    _, nbytes = cs.bvh_blob()
    t = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{cs.device}")
    cs.bvh_export_blob(t, nbytes)
    cs.sync()
    words = t.cpu().numpy().view(np.uint32)
    assert words[1] == n
    n_nodes = max(n - 1, 1)
    q = words[64 : 64 + 8 * n_nodes].reshape(n_nodes, 8)
    lf = words[64 + 8 * n_nodes : 64 + 8 * n_nodes + 8 * n].reshape(n, 8)
    return q, lf


# ---------------------------------------------------------------------------------------------
# golden fixtures (outputs of the real reference)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_golden_scene(CollisionSystem, port, emu, name):
    g = load_golden(name)
    s = g["scene"]
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    # a1: TransformSystem::onUpdate
    assert (bits(cs.world_matrices()) == bits(g["world"])).all(), "world matrices must be bit-exact"
    # a2: refit
    boxes = cs.world_aabbs()
    assert (bits(boxes) == bits(g["world_aabb"])).all(), "world AABBs must be bit-exact"
    # a6/a7 reference-topology mode on the reference's own tree: everything incl. tie winners
    cs.upload_ref_bvh(g["bvh"])
    assert_raycast_equal(cs.get_raycasts(g["rays"], capi.MODE_REFTOPO), g["raycast"], name + " reftopo")
    # a4: the LBVH equals the host execution of the same algorithm bit for bit (sort order, nodes, root)
    order = cs.leaf_order()
    tree = emu.build(boxes, s.collider_entity)
    assert (order == tree["sorted_prim"]).all(), "radix sort / Morton keys disagree with the host restatement"
    exported = cs.bvh
    assert (exported.view(np.uint8) == emu.export(tree).view(np.uint8)).all(), "exported tree differs"
    # a6/a7 production traversal
    got = cs.get_raycasts(g["rays"])
    brute, ties = port.raycast_brute(s.collider_entity[order], boxes[order], g["rays"], nthreads=4, want_ties=True)
    assert_raycast_equal(got, brute, name + " lbvh vs brute force (last-in-Morton-order tie rule)")
    assert_raycast_equal(got, port.raycast_tree(exported, g["rays"]), name + " lbvh vs reference traversal of the exported tree")
    assert_raycast_equal(got, g["raycast"], name + " lbvh vs reference on its own tree", ignore_entity_on=ties > 0)
    if RefEngine.available():
        r = RefEngine()
        r.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
        r.add_colliders(s.collider_entity, s.collider_aabb)
        r.set_bvh(exported)
        assert_raycast_equal(got, r.raycast(g["rays"]), name + " lbvh vs LIVE reference GetRaycast on the exported tree")
    cs.close()


def test_world_matrices_supplied_by_caller(CollisionSystem):
    """The shim path that keeps the reference's CPU TransformSystem: upload world_matrix, skip propagation."""
    g = load_golden("hier1500")
    s = g["scene"]
    cs = CollisionSystem()
    world = np.zeros((s.n_entities + 1, 16), np.float32)
    world[1:] = g["world"]
    cs.set_world_matrices(world)
    cs.set_colliders(s.collider_entity, s.collider_aabb)
    cs.on_update()
    assert (bits(cs.world_aabbs()) == bits(g["world_aabb"])).all()


# ---------------------------------------------------------------------------------------------
# edge cases
# ---------------------------------------------------------------------------------------------
def test_queries_before_first_update_miss(CollisionSystem):
    cs = CollisionSystem()  # frame 0: bvh_ empty, every query misses (reference :164-167)
    out = cs.get_raycasts(synth.random_rays(100, extent=5))
    assert not out.view(np.uint8).any()
    assert cs.get_raycasts(np.zeros((0, 6), np.float32)).shape == (0,)


def test_zero_colliders(CollisionSystem):
    cs = CollisionSystem()
    s = synth.uniform_boxes(4)
    cs.set_hierarchy(s.parent)
    cs.set_trs(s.pos, s.rot, s.scl)
    cs.set_colliders(np.zeros(0, np.uint32), np.zeros((0, 6), np.float32))
    cs.on_update()  # the reference abort()s here (:136 -> :208); we build an empty tree
    assert not cs.get_raycasts(synth.random_rays(64, extent=5)).view(np.uint8).any()
    assert cs.bvh.shape == (0,)


def test_error_statuses(CollisionSystem):
    cs = CollisionSystem()
    with pytest.raises(capi.B2ColError) as e:
        cs.on_update()
    assert e.value.status == capi.ERR_STATE
    with pytest.raises(capi.B2ColError) as e:
        cs.set_hierarchy(np.array([0, 2, 0], np.uint32))  # entity 1's parent is 2
    assert e.value.status == capi.ERR_UNSUPPORTED
    with pytest.raises(capi.B2ColError) as e:
        cs.set_colliders(np.array([0], np.uint32), np.zeros((1, 6), np.float32))
    assert e.value.status == capi.ERR_INVALID
    bad = np.zeros(1, capi.NODE_DTYPE)
    bad["index1"] = 7  # BoundingVolume child outside the array
    with pytest.raises(capi.B2ColError):
        cs.upload_ref_bvh(bad)


def test_degenerate_inputs(CollisionSystem, port):
    """inf / NaN components and zero-length directions must not hang or fault, and must agree with the
    oracle on hit flags and IDs."""
    s = synth.uniform_boxes(2000, seed=51, extent=30)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    rays = synth.random_rays(512, seed=52, extent=35)
    rays[0, 3:] = 0.0                      # zero direction -> NaN after normalize
    rays[1, 3:] = (np.inf, 0, 0)
    rays[2, 0] = np.nan
    rays[3, 3:] = (0.0, -0.0, 1.0)
    rays[4, 3:] = (1e-30, 1e-30, 1e-30)    # denormal dot product
    order, boxes = cs.leaf_order(), cs.world_aabbs()
    got = cs.get_raycasts(rays)
    want = port.raycast_brute(s.collider_entity[order], boxes[order], rays, nthreads=4)
    assert (got["hit"] == want["hit"]).all() and (got["hit_entity"] == want["hit_entity"]).all()


def nan_eq_bits(a, b):
    """bitwise equality of float arrays except that any NaN matches any NaN (x86 and the GPU produce different NaN payloads)"""
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


@pytest.mark.parametrize("with_nan", [False, True])
def test_degenerate_colliders(CollisionSystem, port, emu, with_nan):
    """NaN / +-inf positions, zero / 1e30 / infinite / negative scales, NaN and zero quaternions, a local box whose min exceeds its
    max, an infinite local box — mixed into a 5k scene (reference transformBox src/system_collisions.cpp:14-41 and
    slab test :62-100, whose fminf / fmaxf give these inputs defined results).  World matrices and world AABBs equal the oracle's
    bit for bit (any NaN == any NaN); the tree equals the host restatement incl. its WILD_BIT slots; every ray — sane and degenerate
    ones — gets the reference's leaf-level answer (brute force with the reference's slab test), and, where no NaN plane makes the
    reference's own node unions order-dependent, the answer of the LIVE reference walking the exported tree."""
    s = synth.degenerate_scene(5000, with_nan=with_nan)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    world = port.transform_update(s.parent, s.pos, s.rot, s.scl)
    assert nan_eq_bits(cs.world_matrices(), world[1:]).all(), "world matrices"
    want_boxes = port.refit(s.collider_entity, s.collider_aabb, world).view(np.float32).reshape(-1, 6)
    boxes = cs.world_aabbs()
    assert nan_eq_bits(boxes, want_boxes).all(), "world AABBs"
    assert (~np.isfinite(boxes)).any(axis=1).sum() >= 10
    tree = emu.build(boxes, s.collider_entity)
    assert tree["has_wild"] == 1
    order = cs.leaf_order()
    assert np.array_equal(order, tree["sorted_prim"])
    qnodes, leaves = read_blob(cs, s.n_colliders)
    assert np.array_equal(qnodes, tree["qnodes"]["w"]), "quantised nodes (incl. WILD_BIT flags) differ from the host restatement"
    assert (qnodes[:, 6:8] & 0x40000000).any()
    rays = np.concatenate([synth.degenerate_rays(6000), synth.random_rays(20000, seed=5, extent=70), synth.axis_rays(2000, seed=6, extent=60)])
    got = cs.get_raycasts(rays)
    brute = port.raycast_brute(s.collider_entity[order], boxes[order], rays, nthreads=8)
    assert_raycast_equal(got, brute, "degenerate colliders vs brute force", nan_equal=True)
    # the same scene without the boxes that swallow every ray at t = -inf (infinite / 1e30 scale), so that the other kinds decide
    with np.errstate(invalid="ignore"):
        keep = ~((boxes[:, 3:] - boxes[:, :3]) > 1e6).all(axis=1)
    cs2 = CollisionSystem()
    cs2.set_hierarchy(s.parent)
    cs2.set_trs(s.pos, s.rot, s.scl)
    cs2.set_colliders(s.collider_entity[keep], s.collider_aabb[keep])
    cs2.on_update()
    order2, boxes2 = cs2.leaf_order(), cs2.world_aabbs()
    got2 = cs2.get_raycasts(rays)
    assert_raycast_equal(got2, port.raycast_brute(s.collider_entity[keep][order2], boxes2[order2], rays, nthreads=8), "degenerate colliders (2) vs brute force", nan_equal=True)
    if not with_nan:
        assert 100 < got2["hit"].sum() < rays.shape[0] - 100
        assert_raycast_equal(got2, reference_walk(cs2.bvh, rays, port), "degenerate colliders vs the reference's GetRaycast on the exported tree", nan_equal=True)
        assert_raycast_equal(got, reference_walk(cs.bvh, rays, port), "degenerate colliders (all) vs the reference's GetRaycast on the exported tree", nan_equal=True)
    # refit-only frame keeps the flags
    cs.on_update_refit()
    assert_raycast_equal(cs.get_raycasts(rays), got, "refit frame", nan_equal=True)
    # later builds of the same collider set: adapted key width, keys made inside the refit kernel from a plan based on the bounds the
    # earlier frames reported (non-finite centroids stay out of those bounds and clamp into the plan's edge cells): same answers,
    # and the tree is the host restatement's for that width and plan
    for frame in range(4):
        cs.on_update()
    st = cs.stats()
    assert st["fused_keys"] == 1
    assert nan_eq_bits(cs.world_aabbs(), want_boxes).all(), "world AABBs of the fused-keys frame"
    tree = emu.build(boxes, s.collider_entity, key_bits=st["key_bits"], plan_bounds=(st["scene_min"], st["scene_max"]))
    assert tree["has_wild"] == 1 and np.array_equal(cs.leaf_order(), tree["sorted_prim"])
    qnodes, _ = read_blob(cs, s.n_colliders)
    assert np.array_equal(qnodes, tree["qnodes"]["w"]), "fused-keys frame: quantised nodes differ from the host restatement"
    order = cs.leaf_order()
    assert_raycast_equal(cs.get_raycasts(rays), port.raycast_brute(s.collider_entity[order], boxes[order], rays, nthreads=8), "fused-keys frame vs brute force", nan_equal=True)


def test_degenerate_rays_vs_live_reference(CollisionSystem, port):
    """Degenerate RAYS on a sane scene, every one against the reference: zero / denormal / infinite / NaN directions, NaN origin
    coordinates (the reference's fminf / fmaxf then IGNORE that axis and the ray can hit), infinite and far-away origins."""
    s = synth.uniform_boxes(3000, seed=51, extent=40, hmin=0.5, hmax=3.0)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    rays = synth.degenerate_rays(20000, seed=93, extent=45)
    got = cs.get_raycasts(rays)
    order, boxes = cs.leaf_order(), cs.world_aabbs()
    assert_raycast_equal(got, port.raycast_brute(s.collider_entity[order], boxes[order], rays, nthreads=8), "degenerate rays vs brute force", nan_equal=True)
    assert_raycast_equal(got, reference_walk(cs.bvh, rays, port), "degenerate rays vs the reference's GetRaycast on the exported tree", nan_equal=True)
    nan_origin = np.isnan(rays[:, :3]).any(axis=1) & ~np.isnan(rays[:, :3]).all(axis=1) & np.isfinite(rays[:, 3:]).all(axis=1)
    assert got["hit"][nan_origin].sum() > 10, "rays with one NaN origin coordinate hit in the reference (that axis drops out)"


def test_failed_update_keeps_a_consistent_tree(CollisionSystem, port):
    """b2col_update must not leave a half-described tree behind when it fails.  Good update, then a collider set that
    refers to an entity without a transform: the update fails with B2COL_ERR_INVALID and queries keep answering from the old tree."""
    s = synth.uniform_boxes(20000, seed=77, extent=80)
    rays = synth.random_rays(5000, seed=78, extent=90)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    before = cs.get_raycasts(rays)
    assert before["hit"].sum() > 100
    bad_entity = np.concatenate([s.collider_entity, np.array([s.n_entities + 5], np.uint32)])
    cs.set_colliders(bad_entity, np.concatenate([s.collider_aabb, s.collider_aabb[:1]]))
    with pytest.raises(capi.B2ColError) as ei:
        cs.on_update()
    assert ei.value.status == capi.ERR_INVALID
    assert (cs.get_raycasts(rays).view(np.uint8) == before.view(np.uint8)).all(), "the previous tree must keep answering"
    with pytest.raises(capi.B2ColError):
        cs.on_update_refit()  # the collider set changed: no topology to refit
    with pytest.raises(capi.B2ColError):
        cs.bvh  # export needs a tree of the CURRENT collider set
    cs.set_colliders(s.collider_entity, s.collider_aabb)
    cs.on_update()
    assert (cs.get_raycasts(rays).view(np.uint8) == before.view(np.uint8)).all()
    # a malformed reference tree with a cycle is rejected instead of hanging k_trace_reftopo
    bad = np.zeros(3, capi.NODE_DTYPE)
    bad["index1"] = [1, 2, 0]  # 2 -> 0 -> 1 -> 2
    bad["type2"] = 2
    with pytest.raises(capi.B2ColError):
        cs.upload_ref_bvh(bad)


# ---------------------------------------------------------------------------------------------
# radix sort and hierarchy at scale, duplicates
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,bits_per_axis", [(4095, 16), (4096, 0), (4097, 16), (70001, 10), (300000, 0), (300000, 21), (1 << 20, 0), (3_000_000, 0)])
def test_build_matches_host_restatement(CollisionSystem, emu, n, bits_per_axis):  # 0 = automatic key width
    s = synth.uniform_boxes(n, seed=60 + bits_per_axis, extent=200.0) if n < 200000 else synth.clustered_boxes(n, seed=61, n_clusters=50)
    cs = CollisionSystem(morton_bits_per_axis=bits_per_axis)
    cs.set_scene(s)
    cs.on_update()
    boxes = cs.world_aabbs()
    tree = emu.build(boxes, s.collider_entity, bits=bits_per_axis)
    order = cs.leaf_order()
    assert sorted(order.tolist()) == list(range(n)) if n <= 70001 else np.array_equal(np.sort(order), np.arange(n, dtype=np.uint32))
    assert np.array_equal(order, tree["sorted_prim"]), "sorted order differs from a stable sort of the same keys"
    st = cs.stats()
    assert st["root"] == tree["root"] and st["n_internal_nodes"] == n - 1
    # the traversal structure itself (built in grid space with 128-bit exchanges on the device), bit for bit
    qnodes, leaves = read_blob(cs, n)
    assert np.array_equal(qnodes, tree["qnodes"]["w"]), "quantised nodes differ from the host restatement"
    assert np.array_equal(leaves.view(np.uint32), tree["leaves"].view(np.uint32).reshape(-1, 8)), "leaf records differ"
    if n <= 300000:  # the reference-format export (exact boxes, built on demand by k_build_exact)
        assert (cs.bvh.view(np.uint8) == emu.export(tree).view(np.uint8)).all()


@pytest.mark.parametrize("squash,bits_per_axis", [((1.0, 1.0, 0.02), 0), ((0.03, 1.0, 1.0), 0), ((1.0, 0.0, 0.0), 0), ((1.0, 0.001, 0.2), 14), ((0.0, 0.0, 0.0), 0), ((1.0, 1.0, 0.05), 21)])
def test_build_on_flat_and_degenerate_worlds(CollisionSystem, emu, port, squash, bits_per_axis):
    """Worlds that are thin or flat along some axes: the Morton bits go to the axes that have an extent
    (make_morton_plan); the device keys / sort / tree equal the host restatement bit for bit, and rays agree with brute force."""
    n = 60_000
    s = synth.uniform_boxes(n, seed=66, extent=300.0)
    s.pos[1:] *= np.asarray(squash, np.float32)
    cs = CollisionSystem(morton_bits_per_axis=bits_per_axis)
    cs.set_scene(s)
    cs.on_update()
    boxes = cs.world_aabbs()
    tree = emu.build(boxes, s.collider_entity, bits=bits_per_axis)
    order = cs.leaf_order()
    assert np.array_equal(order, tree["sorted_prim"])
    qnodes, leaves = read_blob(cs, n)
    assert np.array_equal(qnodes, tree["qnodes"]["w"]) and cs.stats()["root"] == tree["root"]
    rays = synth.random_rays(3000, seed=67, extent=300.0)
    rays[:, 0:3] *= np.maximum(np.asarray(squash, np.float32), 0.05)
    got = cs.get_raycasts(rays)
    sel = np.arange(0, 3000, 10)
    assert_raycast_equal(got[sel], port.raycast_brute(s.collider_entity[order], boxes[order], rays[sel], nthreads=8), f"flat world {squash}")
    assert_raycast_equal(got, port.raycast_tree(cs.bvh, rays, nthreads=8), f"flat world {squash} vs reference traversal")
@pytest.mark.gpu
def test_adaptive_key_width(CollisionSystem, port, emu):
    """b2col_update's key-width policy: the first build of a collider set uses the static width (32 bits here); the hierarchy pass
    counts the leaves inside runs of more than four equal keys, at the current width and 8 bits narrower, and a later full build
    drops one radix pass where fewer than 1/32 of the leaves would sit in such runs (uniform scene: none at 24 bits) and adds one
    where more than 1/8 do (clustered scene: 28 % at 32 bits).
    Whatever the width: the tree is the host restatement's tree for that width, and every ray agrees with the live reference
    walking the exported tree."""
    s = synth.uniform_boxes(200_000)
    rays = np.concatenate([synth.random_rays(60_000, seed=5), synth.axis_rays(4_000, seed=6, extent=1000.0)])
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    st = cs.stats()
    assert st["key_bits"] == 32 and st["sort_passes"] == 4 and st["fused_keys"] == 0  # first build: static width, exact Morton plan (k_morton)
    wide = cs.get_raycasts(rays)
    seen = []
    for frame in range(5):  # the statistic of build k is collected by frame k+1 and read by the host before build k+2
        cs.on_update()
        seen.append(cs.stats()["key_bits"])
    assert seen[0] == 32 and seen[-3:] == [24, 24, 24] and sorted(seen, reverse=True) == seen, f"the uniform scene should narrow to 24-bit keys once and stay there: {seen}"
    st = cs.stats()
    # later builds also take their keys from a plan made BEFORE the frame (k_refit_keys: no Morton kernel), whose bounds the header records
    assert st["sort_passes"] == 3 and st["fused_keys"] == 1
    tree = emu.build(cs.world_aabbs(), s.collider_entity, key_bits=24, plan_bounds=(st["scene_min"], st["scene_max"]))
    assert (cs.leaf_order() == tree["sorted_prim"]).all(), "leaf order at 24 bits"
    exported = cs.bvh
    assert (exported.view(np.uint8) == emu.export(tree).view(np.uint8)).all(), "24-bit tree differs from the host restatement"
    got = cs.get_raycasts(rays)
    assert_raycast_equal(got, reference_walk(exported, rays, port), "24-bit keys vs the reference's GetRaycast on the exported tree")
    same = got["hit_entity"] == wide["hit_entity"]
    assert (got["hit"] == wide["hit"]).all() and (got["distance"] == wide["distance"]).all() and same.mean() > 0.999  # the rest: exact ties, other leaf order
    cs.on_update_refit()  # refit-only frames keep the narrowed tree
    assert_raycast_equal(cs.get_raycasts(rays), reference_walk(cs.bvh, rays, port), "refit of the 24-bit tree")
    # a new collider set starts from the static width again; a clustered one widens
    c = synth.clustered_boxes(200_000)
    crays = synth.random_rays(40_000, seed=7, extent=2000.0)
    cs.set_scene(c)
    seen = []
    for frame in range(5):
        cs.on_update()
        seen.append(cs.stats()["key_bits"])
    assert seen[0] == 32 and seen[-3:] == [40, 40, 40] and sorted(seen) == seen, f"the clustered scene (28 % of the leaves in long runs at 32 bits) should widen to 40-bit keys once: {seen}"
    st = cs.stats()
    assert st["fused_keys"] == 1
    tree = emu.build(cs.world_aabbs(), c.collider_entity, key_bits=40, plan_bounds=(st["scene_min"], st["scene_max"]))
    assert (cs.leaf_order() == tree["sorted_prim"]).all(), "leaf order at 40 bits"
    assert_raycast_equal(cs.get_raycasts(crays), reference_walk(cs.bvh, crays, port), "clustered scene")
    cs.close()
def test_morton_plan_follows_the_scene(CollisionSystem, port, emu):
    """The keys of frame k come from a Morton plan made BEFORE frame k, from bounds earlier frames reported (k_refit_keys,
    choose_morton_plan).  A scene that outruns its plan (everything scaled up 8x between two frames: most centroids clamp to the
    plan's edge cells), shrinks far inside it (the plan is re-made one frame later) or drifts away must answer every ray like the
    reference walking the exported tree in EVERY frame — a stale plan may only cost tree quality — and the tree must be the host
    restatement's tree for the plan bounds the header records."""
    s = synth.hierarchy_scene(60_000, 80, extent=300)  # three hierarchy levels: the header is written by the LAST level's last block
    cs = CollisionSystem()
    cs.set_scene(s)
    base = s.pos.copy()
    rng = np.random.default_rng(91)
    rays0 = np.concatenate([synth.random_rays(20_000, seed=92, extent=300), synth.axis_rays(2_000, seed=93, extent=300)])
    plans = []
    #            scale  shift
    script = [(1.0, 0.0), (1.0, 0.0), (1.0, 0.0), (8.0, 0.0), (8.0, 0.0), (8.0, 0.0), (0.05, 0.0), (0.05, 0.0), (0.05, 0.0), (1.0, 5000.0), (1.0, 5000.0), (1.0, 5000.0)]
    for frame, (scale, shift) in enumerate(script):
        s.pos[:] = base
        s.pos[1:81] = base[1:81] * np.float32(scale) + np.float32(shift) + rng.uniform(-0.5, 0.5, (80, 3)).astype(np.float32)  # the roots carry everything below them
        cs.set_trs(s.pos, s.rot, s.scl)
        cs.on_update()
        st = cs.stats()
        plans.append((st["fused_keys"], tuple(st["scene_min"]), tuple(st["scene_max"])))
        rays = rays0.copy()
        rays[:, 0:3] = rays0[:, 0:3] * np.float32(scale if scale >= 1.0 else 1.0) + np.float32(shift)
        boxes = cs.world_aabbs()
        pick = rng.integers(0, boxes.shape[0], 10_000)  # ... and rays aimed at colliders, wherever the scene is now
        target = (boxes[pick, 0:3] + boxes[pick, 3:6]) * np.float32(0.5)
        origin = target + rng.standard_normal((10_000, 3)).astype(np.float32) * np.float32(40.0)
        rays = np.concatenate([rays, np.concatenate([origin, target - origin], axis=1).astype(np.float32)])
        got = cs.get_raycasts(rays)
        exported = cs.bvh
        assert_raycast_equal(got, reference_walk(exported, rays, port), f"frame {frame} (scale {scale}, shift {shift}) vs the reference's GetRaycast on the exported tree")
        if frame in (2, 3, 6, 9, 11):  # settled, first frame after each jump (stale plan), settled again
            tree = emu.build(cs.world_aabbs(), s.collider_entity, key_bits=st["key_bits"], plan_bounds=(st["scene_min"], st["scene_max"]) if st["fused_keys"] else None)
            assert (cs.leaf_order() == tree["sorted_prim"]).all(), f"frame {frame}: leaf order differs from the host restatement under the recorded plan"
            assert (exported.view(np.uint8) == emu.export(tree).view(np.uint8)).all(), f"frame {frame}: exported tree"
        assert got["hit"].sum() > 8000, f"frame {frame}: the aimed rays should hit ({int(got['hit'].sum())})"
    assert plans[0][0] == 0 and all(p[0] == 1 for p in plans[1:]), f"exact plan on the first frame, plan-before-frame afterwards: {[p[0] for p in plans]}"
    assert plans[3][1:] == plans[2][1:], "the first frame after the scene grew still runs on the old plan"
    assert plans[5][1:] != plans[2][1:] and plans[8][1:] != plans[5][1:] and plans[11][1:] != plans[8][1:], "the plan must follow the scene within two frames"
    cs.close()
def test_small_random_scenes_over_many_frames(CollisionSystem, port):
    """The per-frame policies (key width, Morton plan, fused keys, dense / list-driven levels, the header written by the last block of
    the last level's launch) on the scenes where their corner cases live: a handful of colliders, one collider, colliders on a
    subset of the entities in shuffled order (no dense level), colliders only on inner entities (the LAST hierarchy level has none),
    several hierarchy depths.  Every frame of every scene: world AABBs bit-exact vs the oracle and every ray equal to the reference's
    GetRaycast on the exported tree."""
    rng = np.random.default_rng(404)
    for case in range(14):
        n_roots = int(rng.integers(1, 6))
        n_col = [1, 2, 3, 7, 33, 128, 129, 255, 257, 700, 2049, 4097, 9000, 20000][case]
        s = synth.hierarchy_scene(n_col, n_roots, seed=500 + case, extent=60.0, deep_fraction=[0.0, 0.1, 0.5][case % 3])
        keep = np.arange(n_col)
        if case % 4 == 1:  # colliders on a shuffled subset of the entities: list-driven levels, entities without collider
            keep = rng.permutation(n_col)[: max(1, n_col * 2 // 3)]
        elif case % 4 == 2 and n_col > 4:  # only the first (inner) collider entities: the deepest level may hold no collider at all
            keep = np.arange(max(1, n_col // 2))
        s.collider_entity = np.ascontiguousarray(s.collider_entity[keep])
        s.collider_aabb = np.ascontiguousarray(s.collider_aabb[keep])
        rays = np.concatenate([synth.random_rays(3000, seed=600 + case, extent=70.0), synth.axis_rays(300, seed=700 + case, extent=60.0)])
        cs = CollisionSystem(use_cuda_graph=bool(case % 2))
        cs.set_scene(s)
        seen = []
        for frame in range(6):
            if frame:
                synth.perturb_roots(s, n_roots, frame, seed=800 + case)
                if frame == 4:
                    s.pos[1 : n_roots + 1] *= np.float32(3.0)  # the scene outruns the plan of the frames before
                cs.set_trs(s.pos, s.rot, s.scl)
            if frame == 3:
                cs.on_update_refit()
            else:
                cs.on_update()
            world = port.transform_update(s.parent, s.pos, s.rot, s.scl)
            want_boxes = port.refit(s.collider_entity, s.collider_aabb, world).view(np.float32).reshape(-1, 6)
            assert (bits(cs.world_aabbs()) == bits(want_boxes)).all(), f"case {case} frame {frame}: world AABBs"
            assert (bits(cs.world_matrices()) == bits(world[1:])).all(), f"case {case} frame {frame}: world matrices"
            got = cs.get_raycasts(rays)
            assert_raycast_equal(got, reference_walk(cs.bvh, rays, port), f"case {case} ({len(keep)} colliders, {n_roots} roots) frame {frame}")
            st = cs.stats()
            seen.append((st["key_bits"], st["fused_keys"]))
        assert seen[0][1] == 0 and seen[5][1] == 1, f"case {case}: exact plan first, plan-before-frame later: {seen}"
        cs.close()
def test_bound_device_trs_arrays(CollisionSystem, port):
    """b2col_bind_trs_device: the context reads the caller's device TRS arrays in place (no per-frame hand-over copy).  Frames in
    which the arrays are changed on the device between updates: world matrices and world AABBs bit-exact vs the oracle, every ray
    equal to the reference walking the exported tree, and everything bitwise equal to a second context fed through b2col_set_trs.
    Flat (dense, list-free kernel) and three-level scenes, with and without the captured graph."""
    for case, s, roots in ((0, synth.uniform_boxes(50_000, extent=300.0), 0), (1, synth.hierarchy_scene(60_000, 70, extent=300), 70)):
        cs = CollisionSystem(use_cuda_graph=bool(case))
        twin = CollisionSystem()
        cs.set_scene(s)
        twin.set_scene(s)
        d_pos, d_rot, d_scl = (torch.from_numpy(a.copy()).cuda() for a in (s.pos, s.rot, s.scl))
        torch.cuda.synchronize()
        cs.bind_trs_device(d_pos, d_rot, d_scl)
        rays = np.concatenate([synth.random_rays(20_000, seed=61, extent=320), synth.axis_rays(2_000, seed=62, extent=300)])
        rng = np.random.default_rng(63 + case)
        for frame in range(6):
            if frame:
                n_move = roots if roots else s.pos.shape[0] - 1
                s.pos[1 : n_move + 1] += rng.uniform(-1.0, 1.0, (n_move, 3)).astype(np.float32)
                if roots:
                    s.rot[1 : roots + 1] = synth.random_unit_quats(rng, roots)
                d_pos.copy_(torch.from_numpy(s.pos))  # the caller's own device-side update of its arrays (default stream) ...
                d_rot.copy_(torch.from_numpy(s.rot))
                torch.cuda.synchronize()              # ... ordered before the update
                twin.set_trs(s.pos, s.rot, s.scl)
            if frame == 4:
                cs.on_update_refit()
                twin.on_update_refit()
            else:
                cs.on_update()
                twin.on_update()
            world = port.transform_update(s.parent, s.pos, s.rot, s.scl)
            want_boxes = port.refit(s.collider_entity, s.collider_aabb, world).view(np.float32).reshape(-1, 6)
            assert (bits(cs.world_matrices()) == bits(world[1:])).all(), f"case {case} frame {frame}: world matrices"
            assert (bits(cs.world_aabbs()) == bits(want_boxes)).all(), f"case {case} frame {frame}: world AABBs"
            got = cs.get_raycasts(rays)
            assert_raycast_equal(got, reference_walk(cs.bvh, rays, port), f"case {case} frame {frame}: bound arrays vs the reference on the exported tree")
            assert (got.view(np.uint8) == twin.get_raycasts(rays).view(np.uint8)).all(), f"case {case} frame {frame}: bound arrays vs b2col_set_trs"
            assert cs.stats()["key_bits"] == twin.stats()["key_bits"] and cs.stats()["fused_keys"] == twin.stats()["fused_keys"]
        # while bound: no sparse row updates, no host arrays; b2col_propagate alone works and stays in step with the arrays
        with pytest.raises(capi.B2ColError) as ei:
            cs.update_trs_rows(np.array([1], np.uint32), s.pos[1:2], s.rot[1:2], s.scl[1:2])
        assert ei.value.status == capi.ERR_STATE
        assert cs.lib.b2col_bind_trs_device(cs.h, cs.n_entities, capi.ptr(s.pos), capi.ptr(s.rot), capi.ptr(s.scl)) == capi.ERR_INVALID  # host memory
        d_pos[1:] += 0.25
        torch.cuda.synchronize()
        s.pos[1:] += np.float32(0.25)
        cs.propagate()
        assert (bits(cs.world_matrices()) == bits(port.transform_update(s.parent, s.pos, s.rot, s.scl)[1:])).all(), "b2col_propagate on bound arrays"
        # b2col_set_trs takes the context back to its own copy: the caller's arrays are no longer read
        cs.set_trs(s.pos, s.rot, s.scl)
        d_pos.zero_()
        torch.cuda.synchronize()
        cs.on_update()
        twin.set_trs(s.pos, s.rot, s.scl)
        twin.on_update()
        assert (cs.get_raycasts(rays).view(np.uint8) == twin.get_raycasts(rays).view(np.uint8)).all(), "after unbinding"
        cs.close()
        twin.close()


def test_duplicate_boxes_and_ties(CollisionSystem, port):
    s = synth.grid_boxes(12, 12, 12, spacing=3.0)
    # append exact duplicates of the first 300 boxes (new entities at the same place)
    E = s.n_entities
    dup = 300
    s2 = synth.Scene(
        parent=np.concatenate([s.parent, np.zeros(dup, np.uint32)]), pos=np.concatenate([s.pos, s.pos[1 : dup + 1]]),
        rot=np.concatenate([s.rot, s.rot[1 : dup + 1]]), scl=np.concatenate([s.scl, s.scl[1 : dup + 1]]),
        collider_entity=np.arange(1, E + dup + 1, dtype=np.uint32), collider_aabb=np.concatenate([s.collider_aabb, s.collider_aabb[:dup]]),
    )
    cs = CollisionSystem()
    cs.set_scene(s2)
    cs.on_update()
    rays = synth.axis_rays(20000, seed=71, extent=36)
    order, boxes = cs.leaf_order(), cs.world_aabbs()
    got = cs.get_raycasts(rays)
    brute, ties = port.raycast_brute(s2.collider_entity[order], boxes[order], rays, nthreads=8, want_ties=True)
    assert (ties > 0).sum() > 200, "scene must produce exact ties"
    assert_raycast_equal(got, brute, "ties: lbvh vs brute force")
    assert_raycast_equal(got, port.raycast_tree(cs.bvh, rays), "ties: lbvh vs reference traversal of exported tree")


# ---------------------------------------------------------------------------------------------
# configs at (or near) BASELINE sizes: sampled oracle + size-independent properties
# ---------------------------------------------------------------------------------------------
def test_config2_100k_camera_rays(CollisionSystem, port):
    s = synth.static_city(100_000)
    rays = synth.camera_rays(1000, 1000)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    got = cs.get_raycasts(rays)
    exported = cs.bvh
    # the live reference's GetRaycast on the exported tree: every ray
    assert_raycast_equal(got, reference_walk(exported, rays, port), "config 2 vs the reference's GetRaycast on the exported tree")
    # brute force on a sample
    order, boxes = cs.leaf_order(), cs.world_aabbs()
    sel = np.arange(0, rays.shape[0], 997)
    assert_raycast_equal(got[sel], port.raycast_brute(s.collider_entity[order], boxes[order], rays[sel], nthreads=8), "config 2 vs brute force")
    assert got["hit"].sum() > 1000


REF_TREE_100K = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "config2_ref_bvh.npz")


def test_config2_reference_built_tree(CollisionSystem, port):
    """BASELINE config 2 with the reference's OWN tree: its O(N^2) BuildNode (src/system_collisions.cpp:205-376) was run once at
    100 000 colliders (oracle/build_ref_tree.py, half an hour of one core, cached in the git-ignored oracle/_ref/ and shipped with
    the snapshot).  The reference-topology kernel walks that tree and must return the reference's answers on all 1M camera rays bit
    for bit — tie winners included; the production LBVH must agree everywhere except on exact ties."""
    if not os.path.exists(REF_TREE_100K):
        pytest.skip("oracle/_ref/config2_ref_bvh.npz absent: run `python -m oracle.build_ref_tree` where /root/reference exists (about 30 min)")
    from oracle.build_ref_tree import scene_fingerprint

    z = np.load(REF_TREE_100K)
    s = synth.static_city(int(z["n"]))
    assert bytes(z["fingerprint"]).decode() == scene_fingerprint(s), "the cached tree belongs to another scene generator"
    ref_bvh = np.ascontiguousarray(z["bvh"]).view(capi.NODE_DTYPE).reshape(-1)
    rays = synth.camera_rays(1000, 1000)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    assert (bits(cs.world_aabbs()) == bits(z["world_aabb"])).all(), "world AABBs differ from the reference's refit"
    want = reference_walk(ref_bvh, rays, port)  # the reference's GetRaycast on the reference's own tree
    assert want["hit"].sum() > 100_000
    cs.upload_ref_bvh(ref_bvh)
    got_ref = cs.get_raycasts(rays, mode=capi.MODE_REFTOPO)
    assert_raycast_equal(got_ref, want, "config 2, reference topology (k_trace_reftopo) vs the reference")
    assert (got_ref.view(np.uint8) == want.view(np.uint8)).all() or (got_ref["distance"] == want["distance"]).all()
    got = cs.get_raycasts(rays)
    order, boxes = cs.leaf_order(), cs.world_aabbs()
    _, ties = port.raycast_brute(s.collider_entity[order], boxes[order], rays[::50], nthreads=8, want_ties=True)
    differ = got["hit_entity"] != want["hit_entity"]
    assert (got["hit"] == want["hit"]).all() and (got["distance"] == want["distance"]).all(), "LBVH vs reference tree: hit flags / distances"
    assert differ.mean() < 1e-3 and not differ[::50][ties == 0].any(), "LBVH and the reference's tree may only disagree on exact tmin ties"


def test_config3_1m_incoherent(CollisionSystem, port):
    n_rays = 1 << 21
    s = synth.uniform_boxes(1_000_000)
    rays = synth.random_rays(n_rays)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    got = cs.get_raycasts(rays)
    frac = got["hit"].mean()
    assert 0.2 < frac < 0.9, frac
    # (1) idempotence + rebuild determinism
    cs.on_update()
    again = cs.get_raycasts(rays)
    assert (got.view(np.uint8) == again.view(np.uint8)).all()
    # (2) sharded == unsharded (what the multi-GPU path relies on)
    parts = np.concatenate([cs.get_raycasts(rays[lo : lo + n_rays // 8]) for lo in range(0, n_rays, n_rays // 8)])
    assert (parts.view(np.uint8) == got.view(np.uint8)).all()
    # (3) reference traversal of the exported tree on a sample, (4) brute force on a smaller sample
    exported = cs.bvh
    assert_raycast_equal(got, reference_walk(exported, rays, port), "config 3 vs the reference's GetRaycast on the exported tree (all rays)")
    order, boxes = cs.leaf_order(), cs.world_aabbs()
    sel = np.arange(0, n_rays, 8192)
    assert_raycast_equal(got[sel], port.raycast_brute(s.collider_entity[order], boxes[order], rays[sel], nthreads=8), "config 3 vs brute force")
    # (5) geometric self-consistency of every hit: location on the hit box's face named by the normal
    h = got["hit"] > 0
    inv = np.empty(s.n_entities + 1, np.int64)
    inv[s.collider_entity] = np.arange(s.n_colliders)
    b = boxes[inv[got["hit_entity"][h]]]
    loc, nrm = got["location"][h], got["normal"][h]
    axis = np.abs(nrm).argmax(1)
    face = np.where(nrm[np.arange(len(axis)), axis] > 0, b[np.arange(len(axis)), 3 + axis], b[np.arange(len(axis)), axis])
    outside = got["distance"][h] >= 0
    assert np.allclose(loc[np.arange(len(axis)), axis][outside], face[outside], rtol=1e-4, atol=1e-2)


def test_config5_clustered_skewed(CollisionSystem, port):
    """Config 5 at a size the oracle finishes in seconds: Zipf-skewed Gaussian clusters + uniform background (long
    runs of equal Morton prefixes, a deep lopsided tree, wide key), rays aimed at colliders so that most of them hit
    inside a dense cluster, plus random rays.  Reference traversal of the exported tree on a sample, brute force
    over all boxes on a smaller one, shard invariance, and the 40-bit key path (forced: 14 bits per axis)."""
    n, n_rays = 1_500_000, 1 << 20
    s = synth.clustered_boxes(n, seed=5, n_clusters=300)
    rng = np.random.default_rng(55)
    target = s.pos[1 + rng.integers(0, n, n_rays // 2)] + rng.uniform(-0.2, 0.2, (n_rays // 2, 3)).astype(np.float32)
    origin = target + rng.standard_normal((n_rays // 2, 3)).astype(np.float32) * 80.0
    aimed = np.concatenate([origin, target - origin], axis=1).astype(np.float32)  # unnormalised directions
    rays = np.concatenate([aimed, synth.random_rays(n_rays // 2, seed=56, extent=2000.0)])
    results = []
    for bits_per_axis in (0, 14):
        cs = CollisionSystem(morton_bits_per_axis=bits_per_axis)
        cs.set_scene(s)
        cs.on_update()
        got = cs.get_raycasts(rays)
        assert got["hit"][: n_rays // 2].mean() > 0.5, got["hit"][: n_rays // 2].mean()
        parts = np.concatenate([cs.get_raycasts(rays[lo : lo + n_rays // 4]) for lo in range(0, n_rays, n_rays // 4)])
        assert (parts.view(np.uint8) == got.view(np.uint8)).all()
        exported = cs.bvh
        assert_raycast_equal(got, reference_walk(exported, rays, port), f"config 5 ({bits_per_axis}) vs the reference's GetRaycast on the exported tree (all rays)")
        order, boxes = cs.leaf_order(), cs.world_aabbs()
        sel = np.arange(0, n_rays, 4096)
        assert_raycast_equal(got[sel], port.raycast_brute(s.collider_entity[order], boxes[order], rays[sel], nthreads=8), f"config 5 ({bits_per_axis}) vs brute force")
        results.append(got)
    # two different trees over the same boxes: identical answers except on exact ties
    a, b = results
    same = a["hit_entity"] == b["hit_entity"]
    assert (a["hit"] == b["hit"]).all() and (a["distance"] == b["distance"]).all() and same.mean() > 0.999


def test_config5_full_size_10m(CollisionSystem, port):
    """BASELINE config 4 at its full collider count (10M clustered AABBs: 40-bit keys, 5 radix passes, a 640 MB tree with
    the compact copy of its top levels switched on automatically).  Size-independent checks: brute force over all 10M
    boxes on a small sample, the reference's traversal of the exported 10M-leaf tree on a larger one, shard invariance."""
    n = 10_000_000
    s = synth.clustered_boxes(n, seed=5)
    rng = np.random.default_rng(58)
    m = 1 << 17
    target = s.pos[1 + rng.integers(0, n, m)] + rng.uniform(-0.1, 0.1, (m, 3)).astype(np.float32)
    origin = target + rng.standard_normal((m, 3)).astype(np.float32) * 50.0
    rays = np.concatenate([np.concatenate([origin, target - origin], axis=1).astype(np.float32), synth.random_rays(m, seed=59, extent=2000.0)])
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    st = cs.stats()
    assert st["sort_passes"] == 5 and st["n_internal_nodes"] == n - 1
    got = cs.get_raycasts(rays)
    assert got["hit"][:m].mean() > 0.5
    parts = np.concatenate([cs.get_raycasts(rays[lo : lo + m // 2]) for lo in range(0, 2 * m, m // 2)])
    assert (parts.view(np.uint8) == got.view(np.uint8)).all()
    _, nbytes = cs.bvh_blob()
    hdr = torch.empty(256, dtype=torch.uint8, device=f"cuda:{cs.device}")
    blob = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{cs.device}")
    cs.bvh_export_blob(blob, nbytes)
    cs.sync()
    words = blob[:256].cpu().numpy().view(np.uint32)
    del blob, hdr
    assert words[25] == (n - 1) + n and 1000 < words[26] <= 16383, "the traversal must start in the compact copy of the top levels"
    order, boxes = cs.leaf_order(), cs.world_aabbs()
    sel = np.arange(0, 2 * m, 2 * m // 192)
    assert_raycast_equal(got[sel], port.raycast_brute(s.collider_entity[order], boxes[order], rays[sel], nthreads=16), "10M vs brute force")
    assert_raycast_equal(got, reference_walk(cs.bvh, rays, port), "10M vs the reference's GetRaycast on the exported tree (all rays)")


def test_compact_top_levels(CollisionSystem, port, monkeypatch):
    """k_top_layout (trees larger than L2; forced here on a small scene): the copy behind the leaf records is the
    breadth-first order of the top levels with refs remapped, the original nodes are untouched, and every ray gets
    the same answer with and without it — after a full build and after a refit-only frame."""
    n, levels = 200_000, 9
    s = synth.clustered_boxes(n, seed=7, n_clusters=40)
    rng = np.random.default_rng(70)
    target = s.pos[1 + rng.integers(0, n, 100_000)]
    origin = target + rng.standard_normal((100_000, 3)).astype(np.float32) * 60.0
    rays = np.concatenate([np.concatenate([origin, target - origin], axis=1).astype(np.float32), synth.random_rays(50_000, seed=71, extent=2000.0),
                           synth.axis_rays(5_000, seed=72, extent=2000.0)])
    monkeypatch.setenv("B2COL_TOP_LEVELS", "0")
    plain = CollisionSystem()
    plain.set_scene(s)
    plain.on_update()
    want = plain.get_raycasts(rays)
    assert want["hit"].sum() > 10_000
    monkeypatch.setenv("B2COL_TOP_LEVELS", str(levels))
    cs = CollisionSystem()
    cs.set_scene(s)
    for frame in range(2):
        if frame == 0:
            cs.on_update()
        else:
            cs.on_update_refit()  # same boxes, same tree: the copy is regenerated
        got = cs.get_raycasts(rays)
        assert (got.view(np.uint8) == want.view(np.uint8)).all()
        _, nbytes = cs.bvh_blob()
        t = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{cs.device}")
        cs.bvh_export_blob(t, nbytes)
        cs.sync()
        words = t.cpu().numpy().view(np.uint32)
        root, trace_root, top_count = int(words[3]), int(words[25]), int(words[26])
        q = words[64 : 64 + 8 * (n - 1)].reshape(n - 1, 8)
        q_plain, _ = read_blob(plain, n)
        assert (q == q_plain).all(), "the original nodes must stay as built"
        top_base = (n - 1) + n
        assert trace_root == top_base
        top = words[64 + 8 * top_base : 64 + 8 * (top_base + top_count)].reshape(top_count, 8)
        # numpy restatement: breadth-first, slot 0 before slot 1
        order, frontier, expect_refs = [], np.array([root], np.int64), []
        for k in range(levels):
            order.append(frontier)
            kids = q[frontier, 6:8]
            internal = (kids & 0x80000000) == 0
            nxt = kids[internal].astype(np.int64)  # row-major: slot 0 then slot 1 of each node, in frontier order
            refs = kids.copy()
            if k + 1 < levels:
                refs[internal] = top_base + sum(len(o) for o in order) + np.arange(nxt.shape[0], dtype=np.uint32)
            expect_refs.append(refs)
            frontier = nxt
            if frontier.size == 0:
                break
        order, expect_refs = np.concatenate(order), np.concatenate(expect_refs)
        assert top_count == order.shape[0]
        assert (top[:, 0:6] == q[order, 0:6]).all()
        assert (top[:, 6:8] == expect_refs).all()
    # a replica that imports the blob traces through the copy as well
    replica = CollisionSystem()
    replica.bvh_import_blob(t, nbytes)
    assert (replica.get_raycasts(rays).view(np.uint8) == want.view(np.uint8)).all()


def test_config4_dynamic_frames(CollisionSystem, port, emu):
    """Per-frame TRS change -> propagate + refit + rebuild; checked against the oracle every frame."""
    n, roots = 200_000, 200
    s = synth.hierarchy_scene(n, roots, extent=500)
    cs = CollisionSystem()
    cs.set_scene(s)
    rays = synth.random_rays(50_000, seed=81, extent=500)
    for frame in range(3):
        synth.perturb_roots(s, roots, frame)
        cs.set_trs(s.pos, s.rot, s.scl)
        cs.on_update()
        world = port.transform_update(s.parent, s.pos, s.rot, s.scl)
        assert (bits(cs.world_matrices()) == bits(world[1:])).all(), f"frame {frame}: world matrices"
        want_boxes = port.refit(s.collider_entity, s.collider_aabb, world).view(np.float32).reshape(-1, 6)
        boxes = cs.world_aabbs()
        assert (bits(boxes) == bits(want_boxes)).all(), f"frame {frame}: world AABBs"
        got = cs.get_raycasts(rays)
        assert_raycast_equal(got, reference_walk(cs.bvh, rays, port), f"frame {frame} vs the reference's GetRaycast on the exported tree (all rays)")


def test_refit_only_frames(CollisionSystem, port):
    """b2col_refit (the reference's TODO at src/system_collisions.cpp:116-120): after one full
    build, frames that only move colliders keep the tree shape and re-bound it.  Every frame: world matrices and
    world AABBs bit-exact, leaf order unchanged, the exported tree is a valid reference-format tree over the NEW
    boxes, and rays agree with the reference traversal of that tree and with brute force."""
    n, roots = 120_000, 150
    s = synth.hierarchy_scene(n, roots, extent=400)
    cs = CollisionSystem()
    cs.set_scene(s)
    with pytest.raises(capi.B2ColError) as ei:
        cs.on_update_refit()  # no tree yet
    assert ei.value.status == capi.ERR_STATE
    cs.on_update()
    order0 = cs.leaf_order()
    rays = np.concatenate([synth.random_rays(40_000, seed=83, extent=400), synth.axis_rays(4_000, seed=84, extent=400)])
    launches = cs.stats()["kernel_launches"]
    for frame in range(1, 4):
        synth.perturb_roots(s, roots, frame)
        cs.set_trs(s.pos, s.rot, s.scl)
        cs.on_update_refit()
        world = port.transform_update(s.parent, s.pos, s.rot, s.scl)
        assert (bits(cs.world_matrices()) == bits(world[1:])).all(), f"frame {frame}: world matrices"
        want_boxes = port.refit(s.collider_entity, s.collider_aabb, world).view(np.float32).reshape(-1, 6)
        boxes = cs.world_aabbs()
        assert (bits(boxes) == bits(want_boxes)).all(), f"frame {frame}: world AABBs"
        assert (cs.leaf_order() == order0).all(), "refit must not reorder leaves"
        got = cs.get_raycasts(rays)
        exported = cs.bvh
        assert_raycast_equal(got, reference_walk(exported, rays, port), f"refit frame {frame} vs the reference's GetRaycast on the exported tree (all rays)")
        bsel = np.arange(0, rays.shape[0], 200)
        brute = port.raycast_brute(s.collider_entity[order0], boxes[order0], rays[bsel], nthreads=8)
        assert_raycast_equal(got[bsel], brute, f"refit frame {frame} vs brute force")
    st = cs.stats()
    assert st["full_builds"] == 1 and st["refits"] == 3
    # 3 refit frames = 3 x (propagate levels + k_refit + k_header + k_build) + 3 x (k_trace + export kernels): far fewer than rebuilds
    assert st["kernel_launches"] > launches
    # a full rebuild afterwards gives the same answers again (different tree, same hits away from ties)
    cs.on_update()
    again = cs.get_raycasts(rays)
    same = again["hit_entity"] == got["hit_entity"]
    assert (again["hit"] == got["hit"]).all() and (again["distance"][same] == got["distance"][same]).all()
    assert same.mean() > 0.999  # the rest are exact ties resolved by a different leaf order
    # changing the collider set invalidates the topology
    cs.set_colliders(s.collider_entity[:-1], s.collider_aabb[:-1])
    with pytest.raises(capi.B2ColError):
        cs.on_update_refit()


def test_device_resident_and_replicated_tree(CollisionSystem):
    """raycast_device on torch CUDA buffers == host path; a second context that imports the BVH blob
    (what each rank does after the NCCL broadcast) answers identically."""
    s = synth.uniform_boxes(50_000, seed=91, extent=300)
    rays = synth.random_rays(200_000, seed=92, extent=300)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    want = cs.get_raycasts(rays)
    d_rays = torch.from_numpy(rays).cuda()
    d_out = torch.empty(rays.shape[0] * 36, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    cs.raycast_device(d_rays, rays.shape[0], d_out)
    cs.sync()
    assert (d_out.cpu().numpy() == want.view(np.uint8)).all()
    p, nbytes = cs.bvh_blob()
    replica = CollisionSystem()
    blob = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    cs.bvh_export_blob(blob, nbytes)
    cs.sync()
    replica.bvh_import_blob(blob, nbytes)
    replica.sync()  # checks the header that travelled behind the import
    assert (replica.get_raycasts(rays).view(np.uint8) == want.view(np.uint8)).all()
    with pytest.raises(capi.B2ColError):
        replica.bvh_import_blob(blob, nbytes - 32)  # not the size of any blob
    blob[:4] = 0x5A  # corrupt the magic: accepted asynchronously, rejected by the next sync, and the context has no tree then
    replica.bvh_import_blob(blob, nbytes)
    with pytest.raises(capi.B2ColError):
        replica.sync()
    assert not replica.get_raycasts(rays[:100]).view(np.uint8).any()
    st0, st1 = cs.stats(), replica.stats()
    assert st0["kernel_launches"] > 0 and st1["kernel_launches"] > 0


@pytest.mark.skipif(not (DropinEngine.available() and RefEngine.available()), reason="drop-in harness not built (needs /root/reference at build time)")
def test_dropin_shim_inside_reference_scene(port):
    """The reference's own Scene::Update loop and GetSystem<CollisionSystem>()->GetRaycast(), once with the
    reference's CollisionSystem and once with engine_b200/shim's replacement compiled in its place."""
    s = synth.game_scene(800)
    rays = np.concatenate([synth.game_rays(40), synth.random_rays(3000, seed=95, extent=120, normalised=False)])
    engines = []
    for cls in (RefEngine, DropinEngine):
        e = cls()
        e.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
        e.add_colliders(s.collider_entity, s.collider_aabb)
        pre = e.raycast(rays[:64])  # frame 0: bvh_ empty, every query misses
        assert not pre.view(np.uint8).any()
        e.scene_update()
        engines.append(e)
    ref, drop = engines
    want = ref.raycast(rays)
    got_single = drop.raycast(rays)        # GetRaycast, one ray per call, as game code issues them
    got_batch = drop.raycast_batch(rays)   # additive batch entry
    assert (got_single.view(np.uint8) == got_batch.view(np.uint8)).all()
    # the shim's public bvh_ (debug overlay) walked by the reference traversal gives the same answers incl. ties
    exported = drop.bvh()
    assert_raycast_equal(got_batch, port.raycast_tree(exported, rays), "shim vs reference traversal of shim.bvh_")
    ent, box = port.tree_leaves_dfs(exported)
    _, ties = port.raycast_brute(ent, box, rays, nthreads=4, want_ties=True)
    assert_raycast_equal(got_batch, want, "shim vs reference CollisionSystem", ignore_entity_on=ties > 0)
    # second frame after moving things
    s.pos[1:20] += 1.5
    for e in engines:
        e.set_trs(np.arange(1, s.n_entities + 1, dtype=np.uint32), s.pos[1:], s.rot[1:], s.scl[1:])
        e.scene_update()
    exported = drop.bvh()
    ent, box = port.tree_leaves_dfs(exported)
    _, ties = port.raycast_brute(ent, box, rays, nthreads=4, want_ties=True)
    assert_raycast_equal(drop.raycast_batch(rays), ref.raycast(rays), "frame 2", ignore_entity_on=ties > 0)


@pytest.mark.skipif(not DropinEngine.available(), reason="drop-in harness not built (needs /root/reference at build time)")
def test_dropin_compact_hits_and_device_sets():
    """Additive API of the drop-in CollisionSystem: GetRaycastHits + ExpandHits give GetRaycasts' bytes; with UseDevices (>= 2 GPUs)
    large batches are cut across the GPUs after one tree broadcast — same bytes again, also after the next update."""
    s = synth.game_scene(5000)
    rays = np.concatenate([synth.game_rays(200), synth.random_rays(600_000, seed=96, extent=150, normalised=False)])
    e = DropinEngine()
    e.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
    e.add_colliders(s.collider_entity, s.collider_aabb)
    e.scene_update()
    want = e.raycast_batch(rays)
    assert want["hit"].sum() > 1000
    hits, expanded = e.raycast_hits_expanded(rays, threads=3)
    assert (expanded.view(np.uint8) == want.view(np.uint8)).all()
    assert ((hits[:, 2] & 1) == want["hit"]).all() and (hits[:, 0] == want["hit_entity"]).all()
    if torch.cuda.device_count() < 2:
        return
    n_dev = min(torch.cuda.device_count(), 8)
    e.shim_option("use_devices", n_dev)
    assert e.shim_option("devices_in_use") == n_dev
    e.shim_option("multi_gpu_min_rays", 100_000)
    assert (e.raycast_batch(rays).view(np.uint8) == want.view(np.uint8)).all()
    s.pos[1:400] += 2.5  # next frame: the replicas must follow the new tree
    e.set_trs(np.arange(1, s.n_entities + 1, dtype=np.uint32), s.pos[1:], s.rot[1:], s.scl[1:])
    e.scene_update()
    multi = e.raycast_batch(rays)
    e.shim_option("multi_gpu_min_rays", 1 << 40)  # back to one GPU: the reference for the new frame
    single = e.raycast_batch(rays)
    assert (multi.view(np.uint8) == single.view(np.uint8)).all()
    assert (single.view(np.uint8) != want.view(np.uint8)).any()


@pytest.mark.skipif(not (Dropin2Engine.available() and RefEngine.available()), reason="drop-in harness not built (needs /root/reference at build time)")
def test_dropin_transform_and_collision_shims_inside_reference_scene(port):
    """the reference's Scene::Update with BOTH systems replaced — TransformSystem (AoS
    TransformComponent gather -> b2col_propagate -> world_matrix scatter) and CollisionSystem (refit on the
    matrices that stayed on the device).  world_matrix must equal the reference's bit for bit, raycasts too."""
    s = synth.hierarchy_scene(3000, 40, seed=17, extent=150)
    ids = np.arange(1, s.n_entities + 1, dtype=np.uint32)
    rays = np.concatenate([synth.random_rays(4000, seed=96, extent=170, normalised=False), synth.axis_rays(500, seed=97, extent=150)])
    engines = []
    for cls in (RefEngine, Dropin2Engine):
        e = cls()
        e.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
        e.add_colliders(s.collider_entity, s.collider_aabb)
        e.scene_update()
        engines.append(e)
    ref, drop = engines
    for frame in range(3):
        assert (bits(drop.world_matrices(ids)) == bits(ref.world_matrices(ids))).all(), f"world_matrix differs in frame {frame}"
        exported = drop.bvh()
        ent, box = port.tree_leaves_dfs(exported)
        _, ties = port.raycast_brute(ent, box, rays, nthreads=4, want_ties=True)
        assert_raycast_equal(drop.raycast_batch(rays), ref.raycast(rays), f"frame {frame}", ignore_entity_on=ties > 0)
        synth.perturb_roots(s, 40, frame + 1)
        for e in engines:
            e.set_trs(ids, s.pos[1:], s.rot[1:], s.scl[1:])
            e.scene_update()


@pytest.mark.skipif(not (Dropin2Engine.available() and RefEngine.available()), reason="drop-in harness not built (needs /root/reference at build time)")
def test_dropin_sparse_updates_static_and_moving_entities(port):
    """The drop-in pair at work on a mostly static scene (dirty tracking through is_static): after the first
    frame only the non-static entities are read and uploaded (b2col_stage/commit_trs_rows) and only they and their descendants get
    world_matrix written back (b2col_fetch_world_rows) — static CHILDREN of moving roots included.  Every frame: world_matrix of
    every entity bit-equal to the unmodified reference's, raycasts equal; entities and colliders added mid-run, a collider box
    edited in place and a dynamic entity turned static are all picked up."""
    s = synth.hierarchy_scene(6000, 60, seed=19, extent=200)
    E = s.n_entities
    ids = np.arange(1, E + 1, dtype=np.uint32)
    roots = ids[:60]
    rng = np.random.default_rng(3)
    moving_roots = roots[::3]                                        # 20 roots move; their (static-flagged) children follow
    loose = np.sort(rng.choice(np.arange(61, E + 1), 400, replace=False)).astype(np.uint32)  # plus 400 non-static colliders
    dynamic = np.union1d(moving_roots, loose).astype(np.uint32)
    static = np.setdiff1d(ids, dynamic)
    rays = np.concatenate([synth.random_rays(6000, seed=98, extent=230, normalised=False), synth.axis_rays(500, seed=99, extent=200)])
    engines = []
    for cls in (RefEngine, Dropin2Engine):
        e = cls()
        e.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
        e.add_colliders(s.collider_entity, s.collider_aabb)
        e.set_static(static, True)
        e.scene_update()
        engines.append(e)
    ref, drop = engines

    def check(what, all_ids):
        assert (bits(drop.world_matrices(all_ids)) == bits(ref.world_matrices(all_ids))).all(), f"world_matrix differs: {what}"
        ent, box = port.tree_leaves_dfs(drop.bvh())
        _, ties = port.raycast_brute(ent, box, rays, nthreads=4, want_ties=True)
        assert_raycast_equal(drop.raycast_batch(rays), ref.raycast(rays), what, ignore_entity_on=ties > 0)

    check("first frame", ids)
    pos, rot = s.pos.copy(), s.rot.copy()
    for frame in range(4):
        pos[dynamic] += rng.uniform(-1.0, 1.0, (dynamic.shape[0], 3)).astype(np.float32)
        rot[moving_roots] = synth.random_unit_quats(rng, moving_roots.shape[0])
        for e in engines:
            e.set_trs(dynamic, pos[dynamic], rot[dynamic], s.scl[dynamic])
            e.scene_update()
        check(f"sparse frame {frame}", ids)
    assert drop.shim_option("full_uploads") == 1 and drop.shim_option("sparse_uploads") == 4
    # new entities + colliders mid-run (a full upload), then sparse frames again
    new_parent = np.array([0, int(moving_roots[0]), int(moving_roots[0])], np.uint32)
    new_pos = np.array([[10, 20, 30], [1, 2, 3], [-2, 1, 0]], np.float32)
    new_rot = synth.random_unit_quats(rng, 3)
    new_scl = np.ones((3, 3), np.float32)
    for e in engines:
        new_ids = e.create_entities(new_parent, new_pos, new_rot, new_scl)
        e.add_colliders(new_ids, np.tile(np.array([[-1, -1, -1, 1, 1, 1]], np.float32), (3, 1)))
        e.scene_update()
    ids2 = np.arange(1, E + 4, dtype=np.uint32)
    check("after adding entities", ids2)
    assert drop.shim_option("full_uploads") == 2
    # a dynamic entity turns static: forces the long road once, results stay equal
    for e in engines:
        e.set_static(loose[:1], True)
        e.scene_update()
    check("dynamic entity turned static", ids2)
    pos[dynamic] += np.float32(0.5)
    still = dynamic[dynamic != loose[0]]  # is_static is a promise that the TRS no longer changes: the shim does not re-read such rows
    for e in engines:
        e.set_trs(still, pos[still], rot[still], s.scl[still])
        e.scene_update()
    check("sparse frame after the change", ids2)
    # a collider box edited in place is seen by the per-frame comparison (collider_check_interval = 1, the reference's semantics)
    for e in engines:
        e.add_colliders(s.collider_entity[:5], s.collider_aabb[:5] * np.float32(3.0))  # AddComponent again = overwrite the box
        e.scene_update()
    check("collider boxes edited", ids2)


def _game(cls, scene, cam_pos, batched=False, rebuild_interval=1):
    e = cls()
    e.create_entities(scene.parent[1:], scene.pos[1:], scene.rot[1:], scene.scl[1:])
    e.add_colliders(scene.collider_entity, scene.collider_aabb)
    e.add_camera(cam_pos, batched=batched)
    if rebuild_interval != 1:
        e.set_rebuild_interval(rebuild_interval)
    return e


@pytest.mark.skipif(not (DropinEngine.available() and RefEngine.available()), reason="drop-in harness not built (needs /root/reference at build time)")
def test_game_player_controller_trajectories():
    """the test game's own collision response (UNMODIFIED reference
    test/src/camera_controller.cpp, headless, scripted input) driving the player through the scene for 900 frames:
    (a) inside the all-reference engine, (b) with the drop-in CollisionSystem answering its 21 GetRaycast calls per
    frame, (c) with the batched port (one GetRaycasts call per frame) — and (d) the batched port with refit-only
    frames between rebuilds.  The player's state must be identical in every frame."""
    s = synth.game_scene(140)
    inp = synth.game_inputs(900)
    cam = (8.0, 2.0, 4.0)
    want = _game(RefEngine, s, cam).run_game(inp)
    assert want[:, 12].sum() > 100, "the scripted walk should spend time on the ground"
    single = _game(DropinEngine, s, cam).run_game(inp)
    assert (bits(single) == bits(want)).all(), f"drop-in CollisionSystem under the reference controller: first differing frame {np.argwhere((bits(single) != bits(want)).any(1))[:1]}"
    eb = _game(DropinEngine, s, cam, batched=True)
    batched = eb.run_game(inp)
    assert (bits(batched) == bits(want)).all(), f"batched controller: first differing frame {np.argwhere((bits(batched) != bits(want)).any(1))[:1]}"
    c = eb.batched_counters()
    assert c["frames"] == 900 and c["raycast_calls"] <= 900 + c["recasts"] and c["rays_cast"] <= 21 * 900 + c["recasts"]
    assert c["recasts"] < 450, c  # wall pushes are the exception, so most frames need exactly one call
    er = _game(DropinEngine, s, cam, batched=True, rebuild_interval=8)
    refit = er.run_game(inp)
    assert (bits(refit) == bits(want)).all(), "refit-only frames between rebuilds must not change the trajectory"


@pytest.mark.skipif(not (Dropin2Engine.available() and RefEngine.available()), reason="drop-in harness not built (needs /root/reference at build time)")
def test_game_with_moving_colliders_and_refit_interval(port):
    """Both systems replaced, colliders moving every few frames, rebuild_interval = 4: raycasts after each
    frame still equal the reference's (refit frames keep the tree exact)."""
    s = synth.hierarchy_scene(2500, 30, seed=19, extent=120)
    ids = np.arange(1, s.n_entities + 1, dtype=np.uint32)
    rays = np.concatenate([synth.random_rays(3000, seed=98, extent=140, normalised=False), synth.axis_rays(400, seed=99, extent=120)])
    ref, drop = RefEngine(), Dropin2Engine()
    for e in (ref, drop):
        e.create_entities(s.parent[1:], s.pos[1:], s.rot[1:], s.scl[1:])
        e.add_colliders(s.collider_entity, s.collider_aabb)
    drop.set_rebuild_interval(4)
    for frame in range(6):
        synth.perturb_roots(s, 30, frame + 1)
        for e in (ref, drop):
            e.set_trs(ids, s.pos[1:], s.rot[1:], s.scl[1:])
            e.scene_update()
        assert (bits(drop.world_matrices(ids)) == bits(ref.world_matrices(ids))).all(), f"world_matrix differs in frame {frame}"
        exported = drop.bvh()
        ent, box = port.tree_leaves_dfs(exported)
        _, ties = port.raycast_brute(ent, box, rays, nthreads=4, want_ties=True)
        assert_raycast_equal(drop.raycast_batch(rays), ref.raycast(rays), f"frame {frame}", ignore_entity_on=ties > 0)


def test_gltf_ingested_scene_end_to_end(CollisionSystem, port):
    """.glb -> b2col_gltf (entities, TRS, colliders) -> set_hierarchy / set_trs /
    set_colliders -> update -> raycast; world AABBs and hits against the oracle."""
    from engine_b200 import gltf
    from gltf_synth import make_glb

    s = synth.uniform_boxes(200, seed=61, extent=40)
    roots = [gltf.append_to_scene(s, make_glb(seed, n_nodes=40, n_meshes=5), stem=f"file{seed}") for seed in (21, 22, 23)]
    assert roots[0] == 201 and s.n_colliders > 260
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    world = port.transform_update(s.parent, s.pos, s.rot, s.scl)
    assert (bits(cs.world_matrices()) == bits(world[1:])).all()
    want_boxes = port.refit(s.collider_entity, s.collider_aabb, world).view(np.float32).reshape(-1, 6)
    boxes = cs.world_aabbs()
    assert (bits(boxes) == bits(want_boxes)).all()
    rays = synth.random_rays(20_000, seed=62, extent=45)
    order = cs.leaf_order()
    got = cs.get_raycasts(rays)
    assert got["hit"].sum() > 1000
    assert_raycast_equal(got, port.raycast_brute(s.collider_entity[order], boxes[order], rays, nthreads=8), "glTF scene vs brute force")


def test_counting_mode_matches_and_counts(CollisionSystem, emu):
    """B2COL_MODE_LBVH_COUNT: same bytes out as the production mode; the node-visit total agrees with the host
    execution of the same traversal (tests/emu visit histogram) to within the pruning-order slack (both prune with the
    best hit so far; the kernel defers far children through a stack of entry distances exactly as the host loop does)."""
    import ctypes as C

    import emu_loader

    s = synth.uniform_boxes(30_000, seed=95, extent=150)
    rays = synth.random_rays(20_000, seed=96, extent=150)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    want = cs.get_raycasts(rays)
    c0 = cs.trace_counters()
    got = cs.get_raycasts(rays, capi.MODE_LBVH_COUNT)
    c1 = cs.trace_counters()
    assert (got.view(np.uint8) == want.view(np.uint8)).all()
    assert c1["rays"] - c0["rays"] == rays.shape[0]
    visits = c1["node_visits"] - c0["node_visits"]
    tree = emu.build(cs.world_aabbs(), s.collider_entity)
    hist = np.zeros(128, np.uint64)
    tv, lt = C.c_uint64(0), C.c_uint64(0)
    r6 = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    emu.lib.emu_visit_histogram(emu_loader._p(tree["qnodes"]), emu_loader._p(tree["leaves"]), emu_loader._p(tree["grid"]), C.c_uint32(tree["root"]), emu_loader._p(r6),
                                C.c_size_t(r6.shape[0]), C.c_float(1000.0), emu_loader._p(hist), C.c_int(128), C.byref(tv), C.byref(lt))
    assert visits > rays.shape[0]
    assert abs(visits - tv.value) <= 0.02 * tv.value, (visits, tv.value)


@pytest.mark.parametrize("d2h_mode", ["ce", "sm", "auto"])
def test_host_pipeline_d2h_modes(CollisionSystem, monkeypatch, d2h_mode):
    """The chunked host-buffer pipeline returns the same bytes whether the results travel by the copy engine, by SM
    stores into the page-locked buffer (k_copy_out), or by whichever of the two the context measured to be faster;
    ray counts that are not multiples of the chunk / of four exercise the ragged tail of the copy kernel."""
    monkeypatch.setenv("B2COL_D2H_MODE", d2h_mode)
    s = synth.uniform_boxes(100_000, seed=97, extent=400)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    for R in (1_300_003, 262_145, 1 << 20):
        rays = synth.random_rays(R, seed=98, extent=400)
        h_rays = torch.from_numpy(rays).pin_memory()
        h_out = torch.zeros(R * 36 + 64, dtype=torch.uint8).pin_memory()
        d_rays = h_rays.cuda()
        d_out = torch.empty(R * 36, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        cs.raycast_device(d_rays, R, d_out)
        cs.sync()
        want = d_out.cpu()
        for rep in range(4 if d2h_mode == "auto" else 1):
            h_out.fill_(0xAB)
            cs.raycast_host_pinned(h_rays, R, h_out)
            assert bool((h_out[: R * 36] == want).all()), (d2h_mode, R, rep)
            assert bool((h_out[R * 36 :] == 0xAB).all()), "wrote past the end of the result buffer"
    # pageable host memory takes the copy-engine path whatever the mode
    rays = synth.random_rays(300_000, seed=99, extent=400)
    assert (cs.get_raycasts(rays)["hit"].sum()) > 0


def test_compact_hits_expand_to_the_same_raycasts(CollisionSystem):
    """b2col_raycast_hits_batch / _device + b2col_expand_hits (additive ABI: 12 instead of 36 bytes per ray leave the device):
    the expanded Raycasts equal the 36-byte path's byte for byte, through the chunked host pipeline, the single-shot path and
    from device buffers; misses are all-zero in both forms."""
    s = synth.uniform_boxes(200_000, seed=41, extent=300)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    for n in (1, 31, 5000, 700_000):
        rays = np.concatenate([synth.random_rays(n, seed=42, extent=320, normalised=False), synth.degenerate_rays(min(n, 2000), seed=43, extent=320)])[:max(n, 1)]
        want = cs.get_raycasts(rays)
        hits = cs.get_hits(rays)
        assert ((hits["flags"] & 1) == want["hit"]).all() and (hits["hit_entity"] == want["hit_entity"]).all()
        assert (hits.view(np.uint32).reshape(-1, 3)[want["hit"] == 0] == 0).all()
        got = cs.expand_hits(rays, hits, threads=4)
        assert (got.view(np.uint8) == want.view(np.uint8)).all(), f"expanded hits differ from the Raycast path at n = {n}"
    d_rays = torch.from_numpy(rays).cuda()
    d_hits = torch.empty(rays.shape[0] * 12, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    cs.raycast_hits_device(d_rays, rays.shape[0], d_hits)
    cs.sync()
    assert (d_hits.cpu().numpy() == hits.view(np.uint8)).all()


def test_sparse_trs_rows_and_world_rows(CollisionSystem, port):
    """b2col_update_trs_rows / b2col_fetch_world_rows (the drop-in TransformSystem's sparse frame): updating some rows equals
    re-uploading everything; fetched rows equal the full read-back."""
    s = synth.hierarchy_scene(30_000, 200, seed=23, extent=300)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    rng = np.random.default_rng(4)
    moved = np.sort(rng.choice(np.arange(1, s.n_entities + 1), 3000, replace=False)).astype(np.uint32)
    s.pos[moved] += rng.uniform(-2, 2, (3000, 3)).astype(np.float32)
    s.rot[moved] = synth.random_unit_quats(rng, 3000)
    cs.update_trs_rows(moved, s.pos[moved], s.rot[moved], s.scl[moved])
    cs.on_update()
    world = port.transform_update(s.parent, s.pos, s.rot, s.scl)
    assert (bits(cs.world_matrices()) == bits(world[1:])).all()
    pick = np.sort(rng.choice(np.arange(1, s.n_entities + 1), 5000, replace=False)).astype(np.uint32)
    assert (bits(cs.world_rows(pick)) == bits(world[pick])).all()
    want_boxes = port.refit(s.collider_entity, s.collider_aabb, world).view(np.float32).reshape(-1, 6)
    assert (bits(cs.world_aabbs()) == bits(want_boxes)).all()
    with pytest.raises(capi.B2ColError):
        cs.update_trs_rows(np.array([s.n_entities + 1], np.uint32), np.zeros((1, 3)), np.zeros((1, 4)), np.ones((1, 3)))


def test_few_ray_calls_equal_batch(CollisionSystem):
    """GetRaycast is a batch of one: calls of 1..32 rays take the mapped-staging path (one launch, no copies); they must
    return exactly what the same rays return inside a large batch, across more calls than the counter pool has slots."""
    s = synth.uniform_boxes(30_000, seed=93, extent=120)
    cs = CollisionSystem()
    cs.set_scene(s)
    cs.on_update()
    rays = synth.random_rays(6000, seed=94, extent=130)
    want = cs.get_raycasts(rays)
    launches = cs.stats()["kernel_launches"]
    got = np.zeros_like(want)
    i, k = 0, 0
    while i < len(rays):
        m = 1 if k < 2200 else 1 + (k * 7) % 32
        got[i : i + m] = cs.get_raycasts(rays[i : i + m])
        i += m
        k += 1
    assert k > 2100  # more calls than counter slots: the pool was recycled at least once
    assert (got.view(np.uint8) == want.view(np.uint8)).all()
    assert cs.stats()["kernel_launches"] - launches == k
    one = cs.get_raycast(rays[0, :3], rays[0, 3:])
    assert one.tobytes() == want[0].tobytes()


# ---------------------------------------------------------------------------------------------
# multi-GPU data plane (include/b2col.h "multi-GPU"): needs >= 2 visible GPUs (gpurun --gpus 2)
# ---------------------------------------------------------------------------------------------
needs_2_gpus = pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 visible GPUs (gpurun --gpus N)")


@needs_2_gpus
def test_multi_gpu_one_process(port):
    """One process driving every visible GPU (what engine_b200/shim does): b2col_comm_init_all, replicate the tree with one grouped
    ncclBroadcast, sharded raycast with the gather fused into the trace kernels (peer stores into the destination's window), and
    the host-buffer batch split across the GPUs.  EVERY rank's full slice is compared bitwise with a 1-GPU trace of the same rays."""
    from engine_b200.collision import MultiGpu, shard_bounds

    n_dev = min(torch.cuda.device_count(), 8)
    s = synth.uniform_boxes(400_000, seed=31, extent=500.0)
    n_rays = 3_000_007
    rays = synth.random_rays(n_rays, seed=32, extent=520.0)
    mg = MultiGpu(range(n_dev), use_cuda_graph=True)
    mg.primary.set_scene(s)
    mg.primary.on_update()
    want = mg.primary.get_raycasts(rays)  # 1 GPU, whole batch
    assert want["hit"].sum() > 1000
    mg.replicate(0)
    shards = []
    for r in range(n_dev):
        lo, hi = shard_bounds(n_rays, n_dev, r)
        shards.append(torch.from_numpy(rays[lo:hi]).to(f"cuda:{r}"))
    torch.cuda.synchronize()
    for dst in (0, n_dev - 1):
        addr = mg.raycast_gather(shards, n_rays, dst)
        got = mg.systems[dst].read_raycasts(addr, n_rays)
        assert (got.view(np.uint8) == want.view(np.uint8)).all(), f"gathered on rank {dst}: differs from the 1-GPU trace"
    info = mg.primary.comm_info()
    assert info["n_ranks"] == n_dev and info["single_process"] == 1 and info["gathers"] == 2
    # every replica answers on its own too (the broadcast tree is complete)
    for r in range(1, n_dev):
        lo, hi = shard_bounds(n_rays, n_dev, r)
        assert (mg.systems[r].get_raycasts(rays[lo:hi]).view(np.uint8) == want[lo:hi].view(np.uint8)).all()
    # host buffers: slices uploaded / traced / downloaded by their own GPU, straight into the caller's array
    got = mg.get_raycasts(rays)
    assert (got.view(np.uint8) == want.view(np.uint8)).all()
    small = mg.get_raycasts(rays[:1000])  # too small to split: one GPU
    assert (small.view(np.uint8) == want[:1000].view(np.uint8)).all()
    # a new frame on the primary, replicated again
    s.pos[1:] += np.float32(0.5)
    mg.primary.set_trs(s.pos, s.rot, s.scl)
    mg.primary.on_update()
    want2 = mg.primary.get_raycasts(rays)
    mg.replicate(0)
    addr = mg.raycast_gather(shards, n_rays, 0)
    assert (mg.primary.read_raycasts(addr, n_rays).view(np.uint8) == want2.view(np.uint8)).all()
    mg.close()


@needs_2_gpus
def test_multi_gpu_one_process_per_gpu():
    """One process per GPU (torchrun, the way bench.py runs): tests/mp_comm_worker.py on every visible GPU — unique id over gloo,
    b2col_comm_init_rank, the blob broadcast and the records-broadcast + local rebuild, the fused gather onto rank 0 and onto the last
    rank, each compared bitwise with a 1-GPU trace."""
    import socket
    import subprocess
    import sys

    n_dev = min(torch.cuda.device_count(), 8)
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port_no = sk.getsockname()[1]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n_dev}", "--master-addr", "127.0.0.1", "--master-port", str(port_no),
           os.path.join(root, "tests", "mp_comm_worker.py"), "300000", "2000003"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MP_COMM_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
