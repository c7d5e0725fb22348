#!/bin/bash
# r3y: bound TRS arrays through the bulk-copy kernel when their allocations have room
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bound_device or sparse or config4" 2>&1 | tail -3
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r3y_bench.json 2> gpurun_out/r3y_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r3y_bench.err
python - <<P
import json
d=json.load(open("gpurun_out/r3y_bench.json"))
r=d["rebuild"]; print("rebuild ms", r["ms"], "frame_ms", r["frame_ms"], "bound", r["bound_trs"]["ms"], r["bound_trs"]["frame_ms"])
P
