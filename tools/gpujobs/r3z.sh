#!/bin/bash
# r3z: k_onesweep with its key tile moved by cp.async.bulk (dev variant) vs per-thread loads
B2COL_LIB=variants/libb2col_rsbulk.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "build_matches or adaptive or flat_and_degenerate" 2>&1 | tail -2
for i in 1 2; do
echo "== default"; timeout 300 python tools/time_update.py
echo "== bulk tile load"; B2COL_LIB=variants/libb2col_rsbulk.so timeout 300 python tools/time_update.py
done
N_COL=10000000 timeout 300 python tools/time_update.py 2>/dev/null | head -1
N_COL=10000000 B2COL_LIB=variants/libb2col_rsbulk.so timeout 300 python tools/time_update.py 2>/dev/null | head -1
