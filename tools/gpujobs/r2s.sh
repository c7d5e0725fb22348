#!/bin/bash
# r2s: full GPU suite + bench record of the r2 update pipeline (adaptive key width, fused keys, TMA-fed dense refit, 256-bit stores)
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2s_bench.json 2> gpurun_out/r2s_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2s_bench.err
python - <<P
import json
d=json.load(open("gpurun_out/r2s_bench.json"))
print("value",d["value"],"e2e",d["e2e"]["value"],"compact",d["e2e"]["compact_results"]["value"])
print("rebuild",{k:v for k,v in d["rebuild"].items() if k!="what"})
print("roofline",d["roofline"])
for k in ("config4","config5","update_e2e"):
    if k in d: print(k, json.dumps(d[k])[:900])
P
