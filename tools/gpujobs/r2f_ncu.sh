#!/bin/bash
# ncu evidence for round 2 (one call): launch list of the bench command, --set full of the trace kernel, --set full of one update.
# Every command first runs plainly (exit 0) and then, unchanged, under ncu.
cmd="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras"
$cmd > gpurun_out/r2_plain.json 2> gpurun_out/r2_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches.csv $cmd > gpurun_out/r2_ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_trace_lbvh -s 3 -c 1 -f -o gpurun_out/r2_trace $cmd > gpurun_out/r2_ncu_trace.log 2>&1
echo "trace capture rc=$?"
upd="python tools/run_updates.py"
$upd > gpurun_out/r2_updates_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_propagate|k_refit|k_morton|k_onesweep|k_local_sort|k_build' -s 10 -c 6 -f -o gpurun_out/r2_update $upd > gpurun_out/r2_ncu_update.log 2>&1
echo "update capture rc=$?"
c5="python tools/bench_extras.py config5"
$c5 > gpurun_out/r2_config5_plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:k_trace_lbvh -s 4 -c 7 -f -o gpurun_out/r2_config5_trace $c5 > gpurun_out/r2_ncu_config5.log 2>&1
echo "config5 capture rc=$?"
ls -la gpurun_out | tail -14
