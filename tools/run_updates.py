"""dev tool: a few b2col_update calls on the config-3/4 scene (target of `ncu -k ...` captures)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from engine_b200 import synth  # noqa: E402
from engine_b200.collision import CollisionSystem  # noqa: E402

n = int(os.environ.get("N_COL", 1_000_000))
scene = synth.hierarchy_scene(n, 1000) if os.environ.get("HIER") else synth.uniform_boxes(n)
cs = CollisionSystem()
cs.set_scene(scene)
for _ in range(int(os.environ.get("N_UPDATES", 4))):
    cs.on_update()
print("ok", cs.stats()["kernel_launches"])
