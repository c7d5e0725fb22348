#!/bin/bash
# r2r: 32-byte records stored with one 256-bit store (world matrices, world records, leaf records) vs two 128-bit stores
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or build or refit or golden or config4 or degenerate" 2>&1 | tail -3
for i in 1 2; do
echo "== default (256-bit stores)"; timeout 300 python tools/time_update.py
echo "== 128-bit stores"; B2COL_LIB=variants/libb2col_st128.so timeout 300 python tools/time_update.py
done
echo "== phases default"; B2COL_SORT=lsd timeout 300 python tools/phase_timing.py
echo "== phases 128-bit"; B2COL_LIB=variants/libb2col_st128.so B2COL_SORT=lsd timeout 300 python tools/phase_timing.py
