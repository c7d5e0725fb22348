"""dev tool: phase timeline of the last radix pass (needs a -DB2_SORT_TIMING build, B2COL_LIB=variants/...)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from engine_b200 import capi, synth  # noqa: E402
from engine_b200.collision import CollisionSystem  # noqa: E402

n = int(os.environ.get("N_COL", 1_000_000))
cs = CollisionSystem()
cs.set_scene(synth.uniform_boxes(n))
for _ in range(5):
    cs.on_update()
tiles = (n + 4095) // 4096
buf = np.zeros((tiles, 8), np.uint64)
lib = capi.load()
assert lib.b2col_debug_sort_timing(buf.ctypes.data_as(C.c_void_p), C.c_int(tiles)) == 0
t = buf[:, :6].astype(np.int64)
t0 = t[:, 0].min()
rel = (t - t0) / 1e3
names = ["entry", "loaded", "ranked", "lookback done", "keys written", "end"]
print(f"tiles {tiles}; all times in us relative to the first block's entry")
for k, nm in enumerate(names):
    print(f"  {nm:14s} min {rel[:, k].min():7.2f}  median {np.median(rel[:, k]):7.2f}  max {rel[:, k].max():7.2f}")
d = np.diff(rel, axis=1)
for k in range(5):
    print(f"  phase {names[k]:>13s} -> {names[k + 1]:13s} median {np.median(d[:, k]):6.2f}  max {d[:, k].max():6.2f}")
order = np.argsort(t[:, 0])
print("  entry time of blocks (sorted), every 32nd:", np.round(rel[order, 0][::32], 2).tolist())
