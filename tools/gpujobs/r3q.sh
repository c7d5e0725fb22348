#!/bin/bash
# r3q: k_refit_keys* resident blocks per SM (grid = SMs x this): 4 (default) vs 3 vs 2 vs 1 — fewer blocks flush fewer histogram bins
for b in 2 1; do echo "== $b blocks per SM"; B2COL_LIB=variants/libb2col_rkmb$b.so timeout 300 python tools/time_update.py; done
N_COL=10000000 B2COL_LIB=variants/libb2col_rkmb2.so timeout 300 python tools/time_update.py 2>/dev/null | head -1
N_COL=10000000 timeout 300 python tools/time_update.py 2>/dev/null | head -1
N_COL=100000 B2COL_LIB=variants/libb2col_rkmb2.so timeout 300 python tools/time_update.py 2>/dev/null | head -1
N_COL=100000 timeout 300 python tools/time_update.py 2>/dev/null | head -1
