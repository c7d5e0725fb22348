#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -25
python tools/trace_timing.py | tail -1
python bench.py --steps 5 --warmup 3 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo rc=$?; tail -3 gpurun_out/r2d_bench.err
