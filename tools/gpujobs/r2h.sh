#!/bin/bash
./tools/micro/gather_bench | grep -E "records= *(1048576|2097152)" 
# e2e host-buffer pipeline: one vs two trace streams x largest chunk (rays), full and compact result formats
B2COL_E2E_STREAMS=2 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "host_pipeline or compact_hits or few_ray or config3_1m or multi" 2>&1 | tail -3
for st in 1 2; do for ch in 1048576 2097152 4194304; do
  B2COL_E2E_STREAMS=$st B2COL_E2E_CHUNK=$ch python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2h_${st}_${ch}.json 2> gpurun_out/r2h.err
  python - <<P
import json
d=json.load(open("gpurun_out/r2h_${st}_${ch}.json"))
e=d["e2e"]
print("streams $st chunk $ch: full", round(e["value"],1), "compact", round(e["compact_results"]["value"],1), "match", e["matches_device_path"], e["compact_results"]["with_host_expansion"]["equal_to_full_format"])
P
done; done
