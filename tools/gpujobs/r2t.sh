#!/bin/bash
# r2t: hierarchy pass writes every node with ONE 32-byte store under Karras' index (no child refs published): tests + timings
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for i in 1 2; do echo "== default"; timeout 300 python tools/time_update.py; done
echo "== phases"; B2COL_SORT=lsd timeout 300 python tools/phase_timing.py
echo "== static width, exact plan"; B2COL_ADAPTIVE_KEYS=0 B2COL_FUSE_KEYS=0 timeout 300 python tools/time_update.py
timeout 300 python tools/trace_timing.py | tail -1
N_COL=10000000 timeout 300 python tools/time_update.py 2>&1 | head -1
