#!/bin/bash
# r3j: re-capture of the trace kernel's DRAM traffic for the final build (roofline.traffic is keyed by the hash of the kernel's sources)
cmd="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras"
$cmd > gpurun_out/r3j_plain.json 2> gpurun_out/r3j_plain.err &&
ncu --set full --clock-control none --import-source on -k regex:k_trace_lbvh -s 3 -c 1 -f -o gpurun_out/r2_trace $cmd > gpurun_out/r3j_ncu_trace.log 2>&1
echo "trace capture rc=$?"
c5="python tools/bench_extras.py config5"
$c5 > gpurun_out/r3j_config5_plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:k_trace_lbvh -s 4 -c 7 -f -o gpurun_out/r2_config5_trace $c5 > gpurun_out/r3j_ncu_config5.log 2>&1
echo "config5 capture rc=$?"
ls -la gpurun_out/*.ncu-rep
