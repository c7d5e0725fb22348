#!/bin/bash
# r3l: trace kernel A/B: staging rows out of shared memory / sane-scene instantiation alone
for v in default nowild stage1 stage1_nowild stage1_nowild_b11; do
  echo "== $v"
  if [ $v = default ]; then timeout 300 python tools/trace_timing.py | tail -1; else B2COL_LIB=variants/libb2col_$v.so timeout 300 python tools/trace_timing.py | tail -1; fi
done
