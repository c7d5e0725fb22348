#!/bin/bash
# r3m: N-GPU bench of the final round-2 build (no tests)
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r3m_bench$N.json 2> gpurun_out/r3m_bench$N.err
echo "N=$N rc=$?"; tail -3 gpurun_out/r3m_bench$N.err
python - <<P
import json
d=json.load(open("gpurun_out/r3m_bench$N.json"))
m=d["multi_gpu"]
print("N=$N value", round(d["value"],1), "with_gather", round(d["value_with_gather"],1), "e2e", round(d["e2e"]["value"],1), "compact", d["e2e"].get("compact_results") and round(d["e2e"]["compact_results"]["value"],1))
print("strong", m["strong"]); print("bcast_ms", m["bvh_broadcast_ms"], "records_ms", m["records_broadcast_plus_rebuild_ms"], "bitwise", m["all_ranks_bitwise_equal_to_1gpu_trace"])
print("rebuild", d["rebuild"]["ms"], d["rebuild"]["key_bits"], "clocks", d.get("clocks"))
P
