#!/bin/bash
# MSD sort: correctness first, then A/B timing, then the whole suite and the bench
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "build or refit or degenerate or config" 2>&1 | tail -8
for s in lsd msd; do for b in climb direct warp; do echo "== B2COL_SORT=$s B2COL_BUILD=$b"; B2COL_BUILD=$b B2COL_SORT=$s python tools/time_update.py; done; done
echo "== phases lsd/climb"; B2COL_BUILD=climb B2COL_SORT=lsd python tools/phase_timing.py
echo "== phases msd/warp"; B2COL_BUILD=warp B2COL_SORT=msd python tools/phase_timing.py
echo "== phases msd/direct"; B2COL_BUILD=direct B2COL_SORT=msd python tools/phase_timing.py
python -m pytest tests -m gpu -x -q 2>&1 | tail -25
python tools/trace_timing.py | tail -1
echo "== register top-of-stack variant"; B2COL_LIB=variants/libb2col_regtos.so python tools/trace_timing.py | tail -1
B2COL_LIB=variants/libb2col_regtos.so python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or config3_1m or duplicate" 2>&1 | tail -3
python bench.py --steps 5 --warmup 3 > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; echo rc=$?; tail -3 gpurun_out/r2e_bench.err
