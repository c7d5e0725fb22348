#!/bin/bash
# r2o: k_refit_keys geometry sweep: block 256 / 128, grid in resident waves
for lib in default variants/libb2col_rk128.so; do
  for w in 1 2 4 16; do
    echo "== lib=$lib waves=$w"
    if [ "$lib" = default ]; then B2COL_RK_WAVES=$w python tools/time_update.py; else B2COL_LIB=$lib B2COL_RK_WAVES=$w python tools/time_update.py; fi
  done
done
echo "== phases default"; B2COL_SORT=lsd python tools/phase_timing.py
echo "== phases rk128 waves 2"; B2COL_LIB=variants/libb2col_rk128.so B2COL_RK_WAVES=2 B2COL_SORT=lsd python tools/phase_timing.py
