"""Time b2col_update (propagate + refit + sort + build) on device-resident inputs (dev tool)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from engine_b200 import synth  # noqa: E402
from engine_b200.collision import CollisionSystem  # noqa: E402


def main():
    n = int(os.environ.get("N_COL", 1_000_000))
    for name, scene in ((f"flat {n}", synth.uniform_boxes(n)), (f"hier {n} (1000 roots, depth 3)", synth.hierarchy_scene(n, 1000))):
        cs = CollisionSystem(use_cuda_graph=os.environ.get('GRAPH', '1') != '0')
        cs.set_scene(scene)
        for _ in range(5):
            cs.on_update()  # synchronous: the key-width policy reads the statistic of the previous build
        K = 50
        cs.event_record(0)
        for _ in range(K):
            cs.on_update(wait=False)
        cs.event_record(1)
        upd = cs.event_elapsed_ms(0, 1) / K * 1e3
        for _ in range(5):
            cs.on_update_refit(wait=False)
        cs.event_record(2)
        for _ in range(K):
            cs.on_update_refit(wait=False)
        cs.event_record(3)
        print(f"{name:32s} [{cs.stats()['key_bits']}-bit keys] update {upd:8.1f} us   refit-only {cs.event_elapsed_ms(2, 3) / K * 1e3:8.1f} us", flush=True)
        cs.close()


if __name__ == "__main__":
    main()
