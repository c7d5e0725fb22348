#!/bin/bash
# r3x: k_refit_keys_bulk pipeline depth (tiles in flight per block): 2 (default) / 3 / 4; 3 stages on 3 blocks per SM
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or golden or config4 or plan_follows" 2>&1 | tail -2
echo "== default (2 stages)"; timeout 300 python tools/time_update.py
for v in st3 st4 st3_mb3; do echo "== $v"; B2COL_LIB=variants/libb2col_$v.so timeout 300 python tools/time_update.py; done
echo "== default again"; timeout 300 python tools/time_update.py
