#!/bin/bash
# r2l: adaptive key width + dense refit: GPU suite, update timings A/B, threshold calibration (hier scene), wider keys on clustered scenes
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
echo "== default (adaptive keys, dense refit)"; python tools/time_update.py
echo "== B2COL_ADAPTIVE_KEYS=0"; B2COL_ADAPTIVE_KEYS=0 python tools/time_update.py
echo "== B2COL_ADAPTIVE_KEYS=0 B2COL_NO_DENSE_REFIT=1"; B2COL_ADAPTIVE_KEYS=0 B2COL_NO_DENSE_REFIT=1 python tools/time_update.py
echo "== phases (static width)"; B2COL_ADAPTIVE_KEYS=0 B2COL_SORT=lsd python tools/phase_timing.py
echo "== phases (static width, list-driven refit)"; B2COL_ADAPTIVE_KEYS=0 B2COL_NO_DENSE_REFIT=1 B2COL_SORT=lsd python tools/phase_timing.py
for kb in 24 32; do echo "== hier KEY_BITS=$kb"; B2COL_KEY_BITS=$kb python tools/trace_timing.py --workload hier --colliders 1000000 | tail -1; done
for kb in 32 40 48; do echo "== config5 2M KEY_BITS=$kb"; B2COL_KEY_BITS=$kb python tools/trace_timing.py --workload config5 --colliders 2000000 | tail -1; N_COL=2000000 B2COL_KEY_BITS=$kb python tools/time_update.py 2>&1 | head -1; done
for kb in 40 48 56; do echo "== config5 10M KEY_BITS=$kb"; B2COL_KEY_BITS=$kb python tools/trace_timing.py --workload config5 --colliders 10000000 | tail -1; done
