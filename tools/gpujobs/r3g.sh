#!/bin/bash
# r3g: k_refit_keys_bulk with one TMA pipeline per warp (no block barrier in the loop) vs one per block
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or build or refit or golden or config4 or degenerate or plan_follows or dropin or sparse" 2>&1 | tail -3
for i in 1 2; do
echo "== default (per-warp pipelines)"; timeout 300 python tools/time_update.py
echo "== per-block tiles"; B2COL_LIB=variants/libb2col_rkblock.so timeout 300 python tools/time_update.py
done
echo "== no TMA (k_refit_keys)"; B2COL_TMA=0 timeout 300 python tools/time_update.py
echo "== phases"; B2COL_SORT=lsd timeout 300 python tools/phase_timing.py
