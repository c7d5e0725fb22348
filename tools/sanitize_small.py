"""dev tool: a small scene through every kernel of the steady-state frame (several updates so that the key-width / plan policies
switch to k_refit_keys_bulk + k_refit_keys, a refit-only frame, traces, the no-alloc entry point) — the target of
`compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_small.py`."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from engine_b200 import capi, synth  # noqa: E402
from engine_b200.collision import CollisionSystem  # noqa: E402

n = int(os.environ.get("N_COL", 20000))
for name, s, roots in (("flat", synth.uniform_boxes(n, extent=200.0), 0), ("hier", synth.hierarchy_scene(n, 40, extent=200), 40)):
    cs = CollisionSystem(use_cuda_graph=bool(int(os.environ.get("GRAPH", "0"))))
    cs.set_scene(s)
    rays = synth.random_rays(5000, seed=3, extent=220)
    for frame in range(5):
        if roots:
            synth.perturb_roots(s, roots, frame)
            cs.set_trs(s.pos, s.rot, s.scl)
        cs.on_update()
    st = cs.stats()
    cs.on_update_refit()
    got = cs.get_raycasts(rays)
    out = np.zeros(rays.shape[0], capi.RAYCAST_DTYPE)
    assert cs.lib.b2col_raycast_batch_static(cs.h, capi.ptr(np.ascontiguousarray(rays, np.float32)), rays.shape[0], capi.ptr(out)) == capi.OK
    assert (out["hit"] == got["hit"]).all() and (out["hit_entity"] == got["hit_entity"]).all()
    _ = cs.bvh
    print(name, "ok", "key_bits", st["key_bits"], "fused", st["fused_keys"], "hits", int(got["hit"].sum()), flush=True)
    cs.close()
