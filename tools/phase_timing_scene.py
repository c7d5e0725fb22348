"""dev tool: per-phase times of b2col_update (B2COL_PHASE_EVENTS=1) for one named scene: uniform | clustered | city | hier"""
import ctypes as C
import os
import sys

import numpy as np

os.environ["B2COL_PHASE_EVENTS"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from engine_b200 import capi, synth  # noqa: E402
from engine_b200.collision import CollisionSystem  # noqa: E402

kind, n = sys.argv[1], int(sys.argv[2])
scene = {"uniform": synth.uniform_boxes, "clustered": synth.clustered_boxes, "city": synth.static_city, "hier": lambda m: synth.hierarchy_scene(m, 1000)}[kind](n)
lib = capi.load()
lib.b2col_debug_phase_ms.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
cs = CollisionSystem()
cs.set_scene(scene)
acc = []
for it in range(20):
    cs.on_update()
    buf = np.zeros(24, np.float32)
    k = lib.b2col_debug_phase_ms(cs.h, buf.ctypes.data, 24)
    if it >= 5:
        acc.append(buf[:k].copy())
a = np.median(np.stack(acc), 0) * 1e3
print(os.environ.get("B2COL_LIB", "default"), kind, n, " ".join(f"{v:.1f}" for v in a), f"sum={a.sum():.1f} us", flush=True)
