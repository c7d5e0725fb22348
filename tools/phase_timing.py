"""dev tool: per-phase times of b2col_update from CUDA events between the phases (B2COL_PHASE_EVENTS=1), next to
the un-instrumented total — the difference is launch gaps + event overhead."""
import ctypes as C
import os
import sys

import numpy as np

os.environ["B2COL_PHASE_EVENTS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from engine_b200 import capi, synth  # noqa: E402
from engine_b200.collision import CollisionSystem  # noqa: E402

n = int(os.environ.get("N_COL", 1_000_000))
lib = capi.load()
lib.b2col_debug_phase_ms.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
for name, scene in ((f"flat {n}", synth.uniform_boxes(n)), (f"hier {n}", synth.hierarchy_scene(n, 1000))):
    cs = CollisionSystem()
    cs.set_scene(scene)
    acc = []
    for it in range(30):  # synchronous updates: the key-width / plan policies settle within the first five
        cs.on_update()
        buf = np.zeros(16, np.float32)
        k = lib.b2col_debug_phase_ms(cs.h, buf.ctypes.data, 16)
        if it >= 5:
            acc.append(buf[:k].copy())
    a = np.median(np.stack(acc), 0) * 1e3
    levels = cs.stats()["n_levels"]
    fused = len(a) - cs.stats()["sort_passes"] == 4
    names = ["setup"] + (["propagate+refit(%d launches)" % levels] if fused else ["propagate(%d launches)" % levels, "refit"]) + ["morton"]
    names += [f"radix{p}" for p in range(len(a) - len(names) - 1)] + ["build"]
    st = cs.stats()
    if st["fused_keys"]:
        names = ["setup", "propagate+refit+keys(%d launches)" % levels, "(no morton)"] + [f"radix{p}" for p in range(st["sort_passes"])] + ["build"]
    elif os.environ.get("B2COL_SORT", "") != "lsd":  # MSD pipeline: [setup, transform/refit phases..., morton, msd pass, local sort, build]
        names = [f"p{i}" for i in range(len(a) - 4)] + ["morton", "msd", "local", "build"]
    print(name, f"[{st['key_bits']}-bit keys]", " ".join(f"{nm}={v:.1f}" for nm, v in zip(names, a)), f"sum={a.sum():.1f} us")
    cs.close()
