#!/bin/bash
# r3t: b2col_bind_trs_device: its test, neighbours, then the bench records with the bound-TRS frames
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bound_device or sparse or config4 or adaptive or small_random or dropin" 2>&1 | tail -8
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r3t_bench.json 2> gpurun_out/r3t_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r3t_bench.err
python - <<P
import json
d=json.load(open("gpurun_out/r3t_bench.json"))
r=d["rebuild"]; print("rebuild ms", r["ms"], "frame_ms", r["frame_ms"], "bound", r["bound_trs"]["ms"], r["bound_trs"]["frame_ms"])
c=d["config4"]; print("config4 rebuild", c["rebuild"]); print("config4 bound", {k:v for k,v in c["rebuild_bound_trs"].items() if k!="what"})
P
