#!/bin/bash
# dev: 2-GPU check of the multi-GPU data plane (tests + bench with staged / per-lane remote stores)
python tools/trace_timing.py 2>&1 | tail -1
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_gpu or replicated or device_sets" 2>&1 | tail -15
for st in 1 0; do
  B2COL_REMOTE_STAGE=$st python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2c_bench2_stage$st.json 2> gpurun_out/r2c_bench2.err
  echo rc=$?
  python - <<P
import json
d=json.load(open("gpurun_out/r2c_bench2_stage$st.json"))
print("stage$st", d["value"], d["value_with_gather"], d["e2e"]["value"], d["multi_gpu"]["strong"], d["multi_gpu"]["bvh_broadcast_ms"], d["multi_gpu"]["all_ranks_bitwise_equal_to_1gpu_trace"])
P
done
tail -5 gpurun_out/r2c_bench2.err
