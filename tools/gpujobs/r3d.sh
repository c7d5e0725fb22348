#!/bin/bash
# r3d: REDUX bounds reduction, 32-bit Morton path, plan-follows-scene test
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "plan_follows or adaptive or build or refit or golden or config4 or degenerate or ties" 2>&1 | tail -4
for i in 1 2; do echo "== default"; timeout 300 python tools/time_update.py; done
echo "== phases"; B2COL_SORT=lsd timeout 300 python tools/phase_timing.py
echo "== first-frame pipeline (exact plan, static width)"; B2COL_ADAPTIVE_KEYS=0 B2COL_FUSE_KEYS=0 timeout 300 python tools/time_update.py
