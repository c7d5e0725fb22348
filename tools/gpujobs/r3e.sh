#!/bin/bash
# r3e: k_build block size x compaction rounds (compaction changes the trade-off the r1 block-size sweep measured)
echo "== default (128, 1)"; timeout 300 python tools/time_update.py
for v in bb256_c1 bb256_c2 bb256_c3 bb512_c2 bb512_c3 bb64_c1; do echo "== $v"; B2COL_LIB=variants/libb2col_$v.so timeout 300 python tools/time_update.py; done
