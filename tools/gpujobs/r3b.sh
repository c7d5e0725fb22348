#!/bin/bash
# r3b: final single-GPU validation of round 2: full GPU suite, smoke, bench (with extras and cpu baseline), reference arm
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r3b_bench.json 2> gpurun_out/r3b_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r3b_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r3b_reference.json 2> gpurun_out/r3b_reference.err; echo "reference rc=$?"
python - <<P
import json
d=json.load(open("gpurun_out/r3b_bench.json"))
print("value",d["value"],"e2e",d["e2e"]["value"],"compact",d["e2e"]["compact_results"]["value"], "launches", d.get("gpu_launches"))
print("rebuild",{k:v for k,v in d["rebuild"].items() if k!="what"})
for k in ("config4","config5","update_e2e"):
    if k in d: print(k, json.dumps(d[k])[:1500])
r=json.load(open("gpurun_out/r3b_reference.json")); print("reference", r["value"], r.get("cpu_baseline"))
P
