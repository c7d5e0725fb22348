#!/bin/bash
# r2x: look-back batch / group variants of k_onesweep; default = batch 8, group 8
echo "== default"; timeout 300 python tools/time_update.py
for v in lb4 lb16 lb32 lb16g16; do echo "== $v"; B2COL_LIB=variants/libb2col_$v.so timeout 300 python tools/time_update.py; done
echo "== 10M default / lb32"; N_COL=10000000 timeout 300 python tools/time_update.py | head -1; N_COL=10000000 B2COL_LIB=variants/libb2col_lb32.so timeout 300 python tools/time_update.py | head -1
