#!/bin/bash
# r2q: k_refit_keys_bulk (TMA-fed dense levels): tests, A/B
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or build or refit or golden or config4 or degenerate or dropin or sparse or game" 2>&1 | tail -4
echo "== default (bulk)"; timeout 300 python tools/time_update.py
echo "== B2COL_TMA=0"; B2COL_TMA=0 timeout 300 python tools/time_update.py
echo "== phases default"; B2COL_SORT=lsd timeout 300 python tools/phase_timing.py
echo "== bulk, 2 waves"; B2COL_RK_WAVES=2 timeout 300 python tools/time_update.py
echo "== stride kernel for every frame (bulk where dense)"; B2COL_REFIT=stride timeout 300 python tools/time_update.py
