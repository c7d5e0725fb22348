#!/bin/bash
# r3a: k_build two-phase climb (local through shared memory, then the pending climbers through global exchanges)
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or build or refit or golden or config4 or degenerate or ties or top_levels or config5" 2>&1 | tail -3
for i in 1 2; do echo "== default"; timeout 300 python tools/time_update.py; done
echo "== phases"; B2COL_SORT=lsd timeout 300 python tools/phase_timing.py
N_COL=10000000 timeout 300 python tools/time_update.py 2>/dev/null | head -1
