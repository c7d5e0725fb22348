#!/bin/bash
# 8-GPU confirmation of the multi-GPU data plane: tests on all GPUs, bench at N = 8 (staged / per-lane remote stores) and N = 4
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_gpu or replicated or device_sets" 2>&1 | tail -5
run() {  # n stage tag
  B2COL_REMOTE_STAGE=$2 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 5 --warmup 3 > gpurun_out/r2i_bench$1_$3.json 2> gpurun_out/r2i_bench$1_$3.err
  echo "N=$1 $3 rc=$?"
  python - <<P
import json
try:
    d=json.load(open("gpurun_out/r2i_bench$1_$3.json"))
    m=d["multi_gpu"]
    print("N=$1 $3", "value", round(d["value"],1), "with_gather", round(d["value_with_gather"],1), "e2e", round(d["e2e"]["value"],1), "strong", m["strong"], "bcast_ms", m["bvh_broadcast_ms"], "records_ms", m.get("records_broadcast_plus_rebuild_ms"), "bitwise", m["all_ranks_bitwise_equal_to_1gpu_trace"], m.get("comm"))
except Exception as e:
    print("no record:", e)
P
}
run 8 1 staged
run 8 0 perlane
run 4 1 staged
tail -3 gpurun_out/r2i_bench8_staged.err
