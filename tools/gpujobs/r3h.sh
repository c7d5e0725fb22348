#!/bin/bash
# r3h: where k_refit_keys_bulk's time goes (timing-only variants with wrong results: no matrix store / no key + histograms)
for v in default noworld nokeys nokeys_noworld; do
  echo "== $v"
  if [ $v = default ]; then B2COL_SORT=lsd timeout 300 python tools/phase_timing.py | head -1; else B2COL_LIB=variants/libb2col_$v.so B2COL_SORT=lsd timeout 300 python tools/phase_timing.py | head -1; fi
done
