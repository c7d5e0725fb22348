#!/bin/bash
# r2z: ncu evidence for the round-2 update pipeline: launch list of the bench command + --set full of one steady-state update
# (frame 5 of tools/run_updates.py: key width and Morton plan have settled: k_refit_keys_bulk, k_onesweep x3, k_build).
# Every command first runs plainly (exit 0) and then, unchanged, under ncu.
cmd="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras"
$cmd > gpurun_out/r2z_plain.json 2> gpurun_out/r2z_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2z_launches.csv $cmd > gpurun_out/r2z_ncu_launches.log 2>&1
echo "launch list rc=$?"
upd="env N_UPDATES=8 python tools/run_updates.py"
$upd > gpurun_out/r2z_updates_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_propagate|k_refit|k_morton|k_onesweep|k_local_sort|k_build' -s 23 -c 5 -f -o gpurun_out/r2z_update $upd > gpurun_out/r2z_ncu_update.log 2>&1
echo "update capture rc=$?"
ls -la gpurun_out | tail -8
