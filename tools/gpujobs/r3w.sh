#!/bin/bash
# r3w: refit-only frames through the grid-stride kernel (header by its last block, no k_header) now that it runs on 2 blocks per SM
for i in 1 2; do
echo "== default"; timeout 300 python tools/time_update.py
echo "== B2COL_REFIT=stride"; B2COL_REFIT=stride timeout 300 python tools/time_update.py
done
echo "== stride, no TMA"; B2COL_REFIT=stride B2COL_TMA=0 timeout 300 python tools/time_update.py
