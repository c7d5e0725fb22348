"""Seeded synthetic scenes and ray batches for the BASELINE.json configs.

A scene is what the reference's ECS would hold for the collision path: per-entity TRS + parent
(``TransformComponent``, reference include/component_transform.h:13-25) and, for collider
entities, a local ``AABB`` (include/component_collider.h:7-14).  Arrays are indexed by entity
ID; row 0 is a dummy because entity 0 is invalid (include/scene.h:141).  Quaternions are stored
x,y,z,w (glm memory order).  Parents always have smaller IDs than their children, as in every
reference code path that creates entities (src/scene.cpp:35-53, src/gltf_loader.cpp:603-674).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class Scene:
    parent: np.ndarray  # (E+1,) u32
    pos: np.ndarray  # (E+1,3) f32
    rot: np.ndarray  # (E+1,4) f32 xyzw
    scl: np.ndarray  # (E+1,3) f32
    collider_entity: np.ndarray  # (N,) u32 ascending
    collider_aabb: np.ndarray  # (N,6) f32 local min.xyz,max.xyz

    @property
    def n_entities(self) -> int:
        return self.parent.shape[0] - 1

    @property
    def n_colliders(self) -> int:
        return self.collider_entity.shape[0]


def _empty(E: int) -> Scene:
    rot = np.zeros((E + 1, 4), np.float32)
    rot[:, 3] = 1.0
    return Scene(
        parent=np.zeros(E + 1, np.uint32),
        pos=np.zeros((E + 1, 3), np.float32),
        rot=rot,
        scl=np.ones((E + 1, 3), np.float32),
        collider_entity=np.zeros(0, np.uint32),
        collider_aabb=np.zeros((0, 6), np.float32),
    )


def random_unit_quats(rng: np.random.Generator, n: int) -> np.ndarray:
    q = rng.standard_normal((n, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True).astype(np.float32)
    return q.astype(np.float32)


def uniform_boxes(n: int, seed: int = 3, extent: float = 1000.0, hmin: float = 0.5, hmax: float = 2.0) -> Scene:
    """Config 3 (and the small parity cases): n axis-aligned boxes, centres uniform in
    [-extent, extent]^3, half-extents uniform in [hmin, hmax]; one root entity per box, the box
    centre carried by the entity position (so the refit is exercised), identity rotation."""
    rng = np.random.default_rng(seed)
    s = _empty(n)
    s.pos[1:] = rng.uniform(-extent, extent, (n, 3)).astype(np.float32)
    h = rng.uniform(hmin, hmax, (n, 3)).astype(np.float32)
    s.collider_entity = np.arange(1, n + 1, dtype=np.uint32)
    s.collider_aabb = np.concatenate([-h, h], axis=1).astype(np.float32)
    return s


def static_city(n: int, seed: int = 2) -> Scene:
    """Config 2: centres uniform in [-500,500]^2 x [0,50], half-extents log-uniform [0.25,4],
    identity rotation for half of the boxes and a random unit quaternion for the rest."""
    rng = np.random.default_rng(seed)
    s = _empty(n)
    s.pos[1:, 0:2] = rng.uniform(-500.0, 500.0, (n, 2)).astype(np.float32)
    s.pos[1:, 2] = rng.uniform(0.0, 50.0, n).astype(np.float32)
    h = np.exp(rng.uniform(np.log(0.25), np.log(4.0), (n, 3))).astype(np.float32)
    rotated = rng.random(n) < 0.5
    q = random_unit_quats(rng, n)
    s.rot[1:][rotated] = q[rotated]
    s.collider_entity = np.arange(1, n + 1, dtype=np.uint32)
    s.collider_aabb = np.concatenate([-h, h], axis=1).astype(np.float32)
    return s


def grid_boxes(nx: int, ny: int, nz: int, spacing: float = 4.0, half: float = 1.0) -> Scene:
    """Grid-aligned unit boxes: produces exact tmin ties for axis-parallel rays (the parity hazard
    of SURVEY.md section 0)."""
    n = nx * ny * nz
    s = _empty(n)
    ii, jj, kk = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    s.pos[1:] = (np.stack([ii, jj, kk], -1).reshape(-1, 3) * spacing).astype(np.float32)
    s.collider_entity = np.arange(1, n + 1, dtype=np.uint32)
    s.collider_aabb = np.tile(np.array([-half] * 3 + [half] * 3, np.float32), (n, 1))
    return s


def hierarchy_scene(n_colliders: int, n_roots: int, seed: int = 6, extent: float = 1000.0, deep_fraction: float = 0.1,
                    negative_scale_fraction: float = 0.0) -> Scene:
    """Config 4: n_roots moving root entities (no collider); every collider entity is a child of a
    root (depth 2) and deep_fraction of them hang off another collider instead (depth 3).
    Roots and children carry random rotations and non-uniform scales."""
    rng = np.random.default_rng(seed)
    E = n_roots + n_colliders
    s = _empty(E)
    s.pos[1 : n_roots + 1] = rng.uniform(-extent, extent, (n_roots, 3)).astype(np.float32)
    s.rot[1 : n_roots + 1] = random_unit_quats(rng, n_roots)
    s.scl[1 : n_roots + 1] = rng.uniform(0.5, 2.0, (n_roots, 3)).astype(np.float32)
    n_mid = n_colliders - int(n_colliders * deep_fraction)
    ids = np.arange(n_roots + 1, E + 1, dtype=np.uint32)
    par = np.empty(n_colliders, np.uint32)
    par[:n_mid] = rng.integers(1, n_roots + 1, n_mid, dtype=np.uint32)
    if n_colliders > n_mid:
        par[n_mid:] = rng.integers(n_roots + 1, n_roots + 1 + max(n_mid, 1), n_colliders - n_mid, dtype=np.uint32)
    s.parent[ids] = par
    spread = extent * 0.05
    s.pos[ids] = rng.uniform(-spread, spread, (n_colliders, 3)).astype(np.float32)
    s.rot[ids] = random_unit_quats(rng, n_colliders)
    s.scl[ids] = rng.uniform(0.5, 1.5, (n_colliders, 3)).astype(np.float32)
    if negative_scale_fraction > 0:
        neg = rng.random((n_colliders, 3)) < negative_scale_fraction
        s.scl[ids] = np.where(neg, -s.scl[ids], s.scl[ids]).astype(np.float32)
    h = rng.uniform(0.25, 1.0, (n_colliders, 3)).astype(np.float32)
    c = rng.uniform(-0.5, 0.5, (n_colliders, 3)).astype(np.float32)
    s.collider_entity = ids
    s.collider_aabb = np.concatenate([c - h, c + h], axis=1).astype(np.float32)
    return s


def perturb_roots(s: Scene, n_roots: int, frame: int, seed: int = 7) -> None:
    """Per-frame motion for config 4: seeded translation + re-rotation of the root entities (in place)."""
    rng = np.random.default_rng(seed + frame)
    s.pos[1 : n_roots + 1] += rng.uniform(-1.0, 1.0, (n_roots, 3)).astype(np.float32)
    s.rot[1 : n_roots + 1] = random_unit_quats(rng, n_roots)


def clustered_boxes(n: int, seed: int = 5, n_clusters: int = 1000, sigma: float = 5.0, extent: float = 2000.0,
                    background: float = 0.05) -> Scene:
    """Config 5: Gaussian clusters with Zipf-skewed populations + a uniform background."""
    rng = np.random.default_rng(seed)
    s = _empty(n)
    n_bg = int(n * background)
    n_cl = n - n_bg
    w = 1.0 / np.arange(1, n_clusters + 1, dtype=np.float64)
    w /= w.sum()
    which = rng.choice(n_clusters, size=n_cl, p=w)
    centres = rng.uniform(-extent, extent, (n_clusters, 3))
    p = centres[which] + rng.standard_normal((n_cl, 3)) * sigma
    s.pos[1 : n_cl + 1] = p.astype(np.float32)
    s.pos[n_cl + 1 :] = rng.uniform(-extent, extent, (n_bg, 3)).astype(np.float32)
    h = rng.uniform(0.05, 0.25, (n, 3)).astype(np.float32)
    s.collider_entity = np.arange(1, n + 1, dtype=np.uint32)
    s.collider_aabb = np.concatenate([-h, h], axis=1).astype(np.float32)
    return s


def game_scene(n_target: int = 1000, seed: int = 1) -> Scene:
    """Config 1: a recreation of the demo game's main scene shape (reference test/src/game.cpp):
    16 thin land tiles (local AABB (-eps,-eps,0)-(1+eps,1+eps,0.01) scaled (50,50,5), :203-229), a
    unit cube (:145-149), 80-step staircases under a rotated file-root entity (glTF loader makes
    root -> node -> mesh children, src/gltf_loader.cpp:599-674), then seeded boxes under depth-3
    hierarchies until n_target colliders exist."""
    rng = np.random.default_rng(seed)
    parent, pos, rot, scl, cent, cbox = [0], [[0, 0, 0]], [[0, 0, 0, 1]], [[1, 1, 1]], [], []

    def add(par, p, q=(0, 0, 0, 1), sc=(1, 1, 1), box=None):
        parent.append(par)
        pos.append(list(p))
        rot.append(list(q))
        scl.append(list(sc))
        e = len(parent) - 1
        if box is not None:
            cent.append(e)
            cbox.append(list(box))
        return e

    eps = 0.001
    for i in range(4):
        for j in range(4):
            add(0, (50.0 * (i - 2), 50.0 * (j - 2), 0.0), sc=(50, 50, 5), box=(-eps, -eps, 0, 1 + eps, 1 + eps, 0.01))
    add(0, (10.0, 5.0, 1.0), box=(0, 0, 0, 1, 1, 1))
    r = float(np.sqrt(0.5))
    n_stairs = max(1, min(4, (n_target - 17) // 320))
    for k in range(n_stairs):
        root = add(0, (20.0 * k - 30.0, 15.0, 0.0), q=(r, 0, 0, r))  # Y-up -> Z-up (gltf_loader.cpp:599)
        node = add(root, (0, 0, 0), q=tuple(random_unit_quats(rng, 1)[0] * 0 + np.array([0, 0, 0, 1], np.float32)))
        for stp in range(80):
            add(node, (0.0, 0.25 * stp, 0.4 * stp), box=(-1.0, 0.0, -0.2, 1.0, 0.25, 0.2))
    while len(cent) < n_target:
        root = add(0, rng.uniform(-100, 100, 3) * np.array([1, 1, 0.1]), q=random_unit_quats(rng, 1)[0], sc=rng.uniform(0.5, 3, 3))
        mid = add(root, rng.uniform(-5, 5, 3), q=random_unit_quats(rng, 1)[0], sc=rng.uniform(0.5, 2, 3))
        for _ in range(int(rng.integers(1, 8))):
            if len(cent) >= n_target:
                break
            c = rng.uniform(-2, 2, 3)
            h = rng.uniform(0.2, 1.5, 3)
            add(mid, rng.uniform(-3, 3, 3), q=random_unit_quats(rng, 1)[0], box=tuple(c - h) + tuple(c + h))
    return Scene(
        parent=np.array(parent, np.uint32),
        pos=np.array(pos, np.float32),
        rot=np.array(rot, np.float32),
        scl=np.array(scl, np.float32),
        collider_entity=np.array(cent, np.uint32),
        collider_aabb=np.array(cbox, np.float32).reshape(-1, 6),
    )


# ------------------------------------------------------------------------------------------------
# rays: (n,6) float32 = origin.xyz, direction.xyz (directions NOT normalised on purpose where noted;
# GetRaycast normalises, reference src/system_collisions.cpp:162)
# ------------------------------------------------------------------------------------------------


def random_rays(n: int, seed: int = 4, extent: float = 1000.0, normalised: bool = True) -> np.ndarray:
    """Config 3: origins uniform in the cube, directions uniform on the sphere (i.i.d. => incoherent)."""
    rng = np.random.default_rng(seed)
    out = np.empty((n, 6), np.float32)
    chunk = 1 << 22
    for lo in range(0, n, chunk):
        m = min(chunk, n - lo)
        out[lo : lo + m, 0:3] = rng.uniform(-extent, extent, (m, 3)).astype(np.float32)
        d = rng.standard_normal((m, 3)).astype(np.float32)
        if normalised:
            d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
        out[lo : lo + m, 3:6] = d
    return out


def camera_rays(width: int, height: int, eye=(0.0, 0.0, 100.0), look=(1.0, 0.0, -0.2), vfov_deg: float = 70.0) -> np.ndarray:
    """Config 2: pinhole camera, row-major pixel grid, z-up world (coherent rays)."""
    eye = np.array(eye, np.float64)
    f = np.array(look, np.float64)
    f /= np.linalg.norm(f)
    up = np.array([0.0, 0.0, 1.0])
    rgt = np.cross(f, up)
    rgt /= np.linalg.norm(rgt)
    upv = np.cross(rgt, f)
    th = np.tan(np.radians(vfov_deg) * 0.5)
    aspect = width / height
    xs = (2.0 * (np.arange(width) + 0.5) / width - 1.0) * th * aspect
    ys = (1.0 - 2.0 * (np.arange(height) + 0.5) / height) * th
    X, Y = np.meshgrid(xs, ys, indexing="xy")
    d = f[None, None, :] + X[..., None] * rgt[None, None, :] + Y[..., None] * upv[None, None, :]
    out = np.empty((height * width, 6), np.float32)
    out[:, 0:3] = eye.astype(np.float32)
    out[:, 3:6] = d.reshape(-1, 3).astype(np.float32)
    return out


def axis_rays(n: int, seed: int = 8, extent: float = 20.0, snap: float = 1.0) -> np.ndarray:
    """Axis-parallel and in-plane rays with origins snapped to a lattice: exercises the +-inf
    reciprocal paths and exact tmin ties (the game's fall ray is (0,0,-1), reference
    test/src/camera_controller.cpp:123; its horizontal rays have z = 0, :77-97)."""
    rng = np.random.default_rng(seed)
    o = (np.round(rng.uniform(-extent, extent, (n, 3)) / snap) * snap).astype(np.float32)
    kind = rng.integers(0, 4, n)
    d = np.zeros((n, 3), np.float32)
    axis = rng.integers(0, 3, n)
    sign = rng.choice(np.array([-1.0, 1.0], np.float32), n)
    d[np.arange(n), axis] = sign
    horiz = kind == 1
    ang = rng.uniform(0, 2 * np.pi, n)
    d[horiz] = np.stack([np.cos(ang[horiz]), np.sin(ang[horiz]), np.zeros(horiz.sum())], -1).astype(np.float32)
    diag = kind == 2
    d[diag] = rng.choice(np.array([-1.0, 1.0], np.float32), (int(diag.sum()), 3))
    rnd = kind == 3
    d[rnd] = rng.standard_normal((int(rnd.sum()), 3)).astype(np.float32)
    return np.concatenate([o, d], axis=1).astype(np.float32)


def game_rays(n_frames: int, seed: int = 9) -> np.ndarray:
    """Config 1: per frame the camera controller's 20 coplanar horizontal rays (same direction
    (v.x,v.y,0), origins lerped in z) + 1 vertical fall ray (reference test/src/camera_controller.cpp:77-124)."""
    rng = np.random.default_rng(seed)
    out = np.empty((n_frames, 21, 6), np.float32)
    p = np.array([0.0, 0.0, 3.0])
    for f in range(n_frames):
        p = p + rng.uniform(-0.5, 0.5, 3) * np.array([1, 1, 0.05])
        v = rng.standard_normal(2)
        for k in range(20):
            out[f, k, 0:3] = (p[0], p[1], p[2] - 1.7 + 1.7 * k / 19.0)
            out[f, k, 3:6] = (v[0], v[1], 0.0)
        out[f, 20, 0:3] = p
        out[f, 20, 3:6] = (0.0, 0.0, -1.0)
    return out.reshape(-1, 6)


def game_inputs(n_frames: int, seed: int = 10) -> np.ndarray:
    """Scripted player input for the headless test game (reference test/src/camera_controller.cpp:40-70 reads
    these through the InputManager): rows of (movex, movey, lookx, looky, sprint, jump).  Piecewise-constant
    walking / strafing / turning segments with occasional sprints and jumps, so that the player lands on the land
    tiles, runs into boxes and staircases, climbs, jumps under things and falls off edges."""
    rng = np.random.default_rng(seed)
    out = np.zeros((n_frames, 6), np.float32)
    f = 0
    while f < n_frames:
        seg = int(rng.integers(20, 90))
        movey = float(rng.choice([1.0, 1.0, 1.0, 0.5, 0.0, -1.0]))
        movex = float(rng.choice([0.0, 0.0, 1.0, -1.0, 0.25]))
        turn = float(rng.choice([0.0, 0.0, 6.0, -6.0, 25.0, -25.0]))  # mouse counts per frame (x 0.003 rad)
        look = float(rng.choice([0.0, 0.0, 3.0, -3.0]))
        sprint = float(rng.random() < 0.3)
        out[f : f + seg, 0] = movex
        out[f : f + seg, 1] = movey
        out[f : f + seg, 2] = turn
        out[f : f + seg, 3] = look
        out[f : f + seg, 4] = sprint
        f += seg
    out[:, 5] = (rng.random(n_frames) < 0.02).astype(np.float32)
    return out


def degenerate_scene(n: int = 5000, seed: int = 90, extent: float = 60.0, with_nan: bool = True) -> Scene:
    """A sane uniform scene with degenerate colliders mixed in (the reference's transformBox, src/system_collisions.cpp:14-41,
    and slab test, :62-100, have DEFINED behaviour for them through fminf / fmaxf): infinite and huge positions and scales, zero scale,
    a local box whose min exceeds its max, an infinite local box — and, with_nan, NaN positions / quaternions / box coordinates.
    Every 97th collider starting at 11 is degenerate, cycling through the kinds; the kinds are returned in `scene.degenerate_kind`
    (0 = sane)."""
    rng = np.random.default_rng(seed)
    s = uniform_boxes(n, seed=seed, extent=extent, hmin=0.4, hmax=2.5)
    s.rot[1:] = random_unit_quats(rng, n)  # rotated boxes: the refit matters
    kind = np.zeros(n, np.int32)
    kinds = ["pos_inf", "pos_huge", "scale_zero", "scale_zero_axis", "scale_huge", "scale_inf", "box_inverted", "box_inf", "scale_negative", "pos_1e30"]
    if with_nan:
        kinds += ["pos_nan", "quat_nan", "box_nan_one", "quat_zero"]
    inf, nan = np.float32(np.inf), np.float32(np.nan)
    for j, c in enumerate(range(11, n, 97)):
        k = kinds[j % len(kinds)]
        kind[c] = 1 + j % len(kinds)
        e = c + 1  # entity ID of collider c
        if k == "pos_inf":
            s.pos[e, j % 3] = inf if j % 2 else -inf
        elif k == "pos_huge":
            s.pos[e, j % 3] = np.float32(3.0e38)
        elif k == "pos_1e30":
            s.pos[e] = np.float32(1.0e30)
        elif k == "scale_zero":
            s.scl[e] = 0.0
        elif k == "scale_zero_axis":
            s.scl[e, j % 3] = 0.0
        elif k == "scale_huge":
            s.scl[e] = np.float32(1.0e30)
        elif k == "scale_inf":
            s.scl[e, j % 3] = inf
        elif k == "scale_negative":
            s.scl[e] = (-1.5, 2.0, -0.5)
        elif k == "box_inverted":
            s.collider_aabb[c] = (1.0, 2.0, 0.5, -1.0, -2.0, -0.5)
        elif k == "box_inf":
            s.rot[e] = (0.0, 0.0, 0.0, 1.0)
            s.collider_aabb[c, 3 + j % 3] = inf
        elif k == "pos_nan":
            s.pos[e, j % 3] = nan
        elif k == "quat_nan":
            s.rot[e, j % 4] = nan
        elif k == "quat_zero":
            s.rot[e] = 0.0
        elif k == "box_nan_one":
            s.rot[e] = (0.0, 0.0, 0.0, 1.0)
            s.collider_aabb[c, j % 6] = nan
    s.degenerate_kind = kind
    return s


def degenerate_rays(n: int, seed: int = 91, extent: float = 70.0) -> np.ndarray:
    """Random rays with the degenerate cases of the reference's ray arithmetic mixed in (GetRaycast normalises, :162; the slab test
    divides by the direction, :65): zero / denormal / infinite / NaN direction components, axis-parallel directions with signed
    zeros, NaN and infinite origin coordinates (a NaN coordinate makes the reference IGNORE that axis), origins far outside the scene."""
    rng = np.random.default_rng(seed)
    rays = random_rays(n, seed=seed, extent=extent, normalised=False)
    inf, nan = np.float32(np.inf), np.float32(np.nan)
    for j, i in enumerate(range(3, n, 13)):
        k = j % 14
        if k == 0:
            rays[i, 3:] = 0.0
        elif k == 1:
            rays[i, 3 + j % 3] = inf
        elif k == 2:
            rays[i, j % 3] = nan
        elif k == 3:
            rays[i, 3:] = (0.0, -0.0, 1.0)
        elif k == 4:
            rays[i, 3:] = (1e-30, 1e-30, 1e-30)
        elif k == 5:
            rays[i, j % 3] = inf
        elif k == 6:
            rays[i, 3 + j % 3] = nan
        elif k == 7:
            rays[i, 0:3] = (nan, nan, 0.5)
        elif k == 8:
            rays[i, 0:3] *= np.float32(1.0e6)
        elif k == 9:
            rays[i, 3:] = (-0.0, 1.0, 0.0)
        elif k == 10:
            rays[i, 0:3] = np.float32(1.0e30)
            rays[i, 3:] = (-1.0, -1.0, -1.0)
        elif k == 11:
            rays[i, 3:] = (1.0, 0.0, 0.0)
            rays[i, 0] = -np.float32(3.0e38)
        elif k == 12:
            rays[i, 3 + j % 3] = -inf
        elif k == 13:
            rays[i, 0:3] = rng.uniform(-extent, extent, 3).astype(np.float32)
            rays[i, 3:] = (0.0, 0.0, -1.0)
    return rays
