#!/bin/bash
# r2n: k_refit_keys (grid-stride refit + fused keys from a plan made before the frame): tests, update timings A/B
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or build or refit or golden or config4 or degenerate or dropin or sparse or game" 2>&1 | tail -4
echo "== default"; python tools/time_update.py
echo "== B2COL_FUSE_KEYS=0"; B2COL_FUSE_KEYS=0 python tools/time_update.py
echo "== B2COL_REFIT=classic"; B2COL_REFIT=classic python tools/time_update.py
echo "== B2COL_FUSE_KEYS=0 B2COL_ADAPTIVE_KEYS=0 (r2m pipeline + stride refit)"; B2COL_FUSE_KEYS=0 B2COL_ADAPTIVE_KEYS=0 python tools/time_update.py
echo "== phases default"; B2COL_SORT=lsd python tools/phase_timing.py
echo "== phases FUSE_KEYS=0"; B2COL_FUSE_KEYS=0 B2COL_SORT=lsd python tools/phase_timing.py
