#!/bin/bash
# r3c: N-GPU validation of the final round-2 build: multi-GPU tests, bench at N ranks
N=${1:-2}
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_gpu or replicated or device_sets" 2>&1 | tail -4
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r3c_bench$N.json 2> gpurun_out/r3c_bench$N.err
echo "N=$N rc=$?"; tail -3 gpurun_out/r3c_bench$N.err
python - <<P
import json
d=json.load(open("gpurun_out/r3c_bench$N.json"))
m=d["multi_gpu"]
print("N=$N value", round(d["value"],1), "with_gather", round(d["value_with_gather"],1), "e2e", round(d["e2e"]["value"],1), "compact", d["e2e"].get("compact_results") and round(d["e2e"]["compact_results"]["value"],1))
print("strong", m["strong"]); print("bcast_ms", m["bvh_broadcast_ms"], "records_ms", m["records_broadcast_plus_rebuild_ms"], "bitwise", m["all_ranks_bitwise_equal_to_1gpu_trace"])
print("rebuild", d["rebuild"]["ms"], d["rebuild"]["key_bits"])
P
