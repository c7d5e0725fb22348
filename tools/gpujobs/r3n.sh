#!/bin/bash
# r3n: bulk gather with the halving tail schedule: tests under B2COL_GATHER=bulk on 2 GPUs, then A/B on N GPUs
N=${1:-2}
if [ "$N" = 2 ]; then B2COL_GATHER=bulk timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_gpu or replicated or device_sets" 2>&1 | tail -3; fi
for tail in 1 0; do
  B2COL_GATHER=bulk B2COL_GATHER_TAIL=$tail timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r3n_bench${N}_tail$tail.json 2> gpurun_out/r3n_bench${N}_tail$tail.err
  echo "N=$N tail=$tail rc=$?"
  python - <<P
import json
d=json.load(open("gpurun_out/r3n_bench${N}_tail$tail.json")); m=d["multi_gpu"]
print("N=$N tail=$tail value", round(d["value"],1), "with_gather", round(d["value_with_gather"],1), "strong", round(m["strong"]["ms_per_step_with_gather"],3), "bitwise", m["all_ranks_bitwise_equal_to_1gpu_trace"], m["strong"]["bitwise_equal_to_1gpu_trace"])
P
done
