#!/bin/bash
# bulk gather (chunked local trace + copy-engine shipping) on N GPUs: tests under B2COL_GATHER=bulk, then the bench
N=${1:-2}
B2COL_GATHER=bulk python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_gpu or replicated or device_sets" 2>&1 | tail -5
for g in bulk fused; do
  [ "$N" = 8 ] && [ "$g" = fused ] && continue
  B2COL_GATHER=$g timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2j_bench${N}_$g.json 2> gpurun_out/r2j_bench${N}_$g.err
  echo "N=$N $g rc=$?"
  python - <<P
import json
try:
    d=json.load(open("gpurun_out/r2j_bench${N}_$g.json"))
    m=d["multi_gpu"]
    print("N=$N $g", "value", round(d["value"],1), "with_gather", round(d["value_with_gather"],1), "e2e", round(d["e2e"]["value"],1), "compact", d["e2e"].get("compact_results") and round(d["e2e"]["compact_results"]["value"],1), "strong", m["strong"], "bcast_ms", m["bvh_broadcast_ms"], "bitwise", m["all_ranks_bitwise_equal_to_1gpu_trace"])
except Exception as e:
    print("no record:", e)
P
done
tail -3 gpurun_out/r2j_bench${N}_bulk.err
