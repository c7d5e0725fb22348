#!/bin/bash
# r3f: k_onesweep: tile-local reorder in shared memory BEFORE the look-back wait
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or build or refit or golden or config4 or degenerate or ties or config5 or plan_follows" 2>&1 | tail -3
for i in 1 2; do
echo "== default (overlap)"; timeout 300 python tools/time_update.py
echo "== no overlap"; B2COL_LIB=variants/libb2col_rs_noov.so timeout 300 python tools/time_update.py
done
echo "== static 32-bit, exact plan: overlap / none"; B2COL_ADAPTIVE_KEYS=0 B2COL_FUSE_KEYS=0 timeout 300 python tools/time_update.py | head -1; B2COL_LIB=variants/libb2col_rs_noov.so B2COL_ADAPTIVE_KEYS=0 B2COL_FUSE_KEYS=0 timeout 300 python tools/time_update.py | head -1
echo "== 10M: overlap / none"; N_COL=10000000 timeout 300 python tools/time_update.py 2>/dev/null | head -1; N_COL=10000000 B2COL_LIB=variants/libb2col_rs_noov.so timeout 300 python tools/time_update.py 2>/dev/null | head -1
