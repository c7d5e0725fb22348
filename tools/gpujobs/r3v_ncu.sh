#!/bin/bash
# r3v: refresh of the steady-state update capture for the final build (k_refit_keys_bulk on 2 blocks per SM, REDUX bounds, one fence)
upd="env N_UPDATES=8 python tools/run_updates.py"
$upd > gpurun_out/r3v_updates_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_propagate|k_refit|k_morton|k_onesweep|k_local_sort|k_build' -s 23 -c 5 -f -o gpurun_out/r3v_update $upd > gpurun_out/r3v_ncu_update.log 2>&1
echo "update capture rc=$?"
