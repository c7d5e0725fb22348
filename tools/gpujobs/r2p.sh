#!/bin/bash
# r2p: cache hints on the frame's streams (world matrices / leaf records evict-first), persisting-L2 window over the world records
echo "== default"; python tools/time_update.py
for v in wcs lcs wcs_lcs; do echo "== variant $v"; B2COL_LIB=variants/libb2col_$v.so python tools/time_update.py; done
echo "== L2 persist (graph)"; B2COL_L2_PERSIST=1 python tools/time_update.py
echo "== L2 persist (no graph)"; B2COL_L2_PERSIST=1 GRAPH=0 python tools/time_update.py
echo "== no graph baseline"; GRAPH=0 python tools/time_update.py
echo "== L2 persist + wcs_lcs"; B2COL_L2_PERSIST=1 B2COL_LIB=variants/libb2col_wcs_lcs.so python tools/time_update.py
echo "== phases wcs_lcs"; B2COL_LIB=variants/libb2col_wcs_lcs.so B2COL_SORT=lsd python tools/phase_timing.py
echo "== trace after update, default / wcs_lcs (leaf records evict-first: does the first trace pay?)"
python tools/trace_timing.py | tail -1
B2COL_LIB=variants/libb2col_wcs_lcs.so python tools/trace_timing.py | tail -1
