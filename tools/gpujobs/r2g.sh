#!/bin/bash
# second single-GPU job of round 2: new block-local sort (rounds sized by the bucket), test suite, variants, bench
B2COL_SORT=auto python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "build or refit or degenerate or config or ties" 2>&1 | tail -4
for s in lsd auto; do echo "== B2COL_SORT=$s"; B2COL_SORT=$s python tools/time_update.py; done
echo "== phases auto(msd)"; B2COL_SORT=auto python tools/phase_timing.py
python -m pytest tests -m gpu -x -q 2>&1 | tail -25
python tools/trace_timing.py | tail -1
echo "== register top-of-stack variant"; B2COL_LIB=variants/libb2col_regtos.so python tools/trace_timing.py | tail -1
B2COL_LIB=variants/libb2col_regtos.so python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or config3_1m or duplicate" 2>&1 | tail -3
for s in lsd auto; do B2COL_SORT=$s python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2g_bench_$s.json 2> gpurun_out/r2g_bench_$s.err; echo "bench $s rc=$?"; done
B2COL_E2E_STREAMS=2 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "host_pipeline or compact_hits or few_ray or config3_1m" 2>&1 | tail -3
B2COL_E2E_STREAMS=2 python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2g_bench_streams2.json 2> gpurun_out/r2g_bench_streams2.err; echo "bench streams2 rc=$?"
python bench.py --steps 5 --warmup 3 > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo rc=$?; tail -3 gpurun_out/r2g_bench.err
