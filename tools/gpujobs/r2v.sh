#!/bin/bash
# r2v: k_build with climber compaction after the first levels (B2_BUILD_COMPACT = 0..4)
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or build or refit or golden or config4 or degenerate or ties or top_levels" 2>&1 | tail -3
for c in 0 1 2 3 4; do
  echo "== compact rounds $c"
  if [ $c = 2 ]; then timeout 300 python tools/time_update.py; else B2COL_LIB=variants/libb2col_compact$c.so timeout 300 python tools/time_update.py; fi
done
echo "== phases (2 rounds)"; B2COL_SORT=lsd timeout 300 python tools/phase_timing.py
N_COL=10000000 timeout 300 python tools/time_update.py 2>&1 | head -1
N_COL=10000000 B2COL_LIB=variants/libb2col_compact0.so timeout 300 python tools/time_update.py 2>&1 | head -1
