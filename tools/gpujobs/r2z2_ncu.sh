#!/bin/bash
# launch list of the WHOLE bench command (rebuild measurement + warm-up + the timed trace steps + e2e)
cmd="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
$cmd > gpurun_out/r2z2_plain.json 2> gpurun_out/r2z2_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r2z2_launches.csv $cmd > gpurun_out/r2z2_ncu_launches.log 2>&1
echo "launch list rc=$?"; grep -c k_trace_lbvh gpurun_out/r2z2_launches.csv; wc -l gpurun_out/r2z2_launches.csv
