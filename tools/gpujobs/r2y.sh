#!/bin/bash
# r2y: full suite; key-width policy at 10M uniform (24-bit keys chosen by the run-length statistic: does the traversal pay?)
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for kb in 24 32 40; do B2COL_KEY_BITS=$kb timeout 300 python tools/trace_timing.py --workload config3 --colliders 10000000 | tail -1; done
for kb in 24 32 40; do B2COL_KEY_BITS=$kb timeout 300 python tools/trace_timing.py --workload config3 --colliders 10000000 --aimed | tail -1; done
for kb in 24 32; do B2COL_KEY_BITS=$kb timeout 300 python tools/trace_timing.py --workload config3 --colliders 4000000 | tail -1; done
