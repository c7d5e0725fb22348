#!/bin/bash
# r2w: k_build climber compaction, root finish out of the loop
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or build or refit or golden or config4 or degenerate or ties or top_levels" 2>&1 | tail -3
for c in 0 1 2 3; do
  echo "== compact rounds $c"
  if [ $c = 2 ]; then timeout 300 python tools/time_update.py; else B2COL_LIB=variants/libb2col_compact$c.so timeout 300 python tools/time_update.py; fi
done
echo "== phases (2 rounds)"; B2COL_SORT=lsd timeout 300 python tools/phase_timing.py
