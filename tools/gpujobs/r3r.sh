#!/bin/bash
# r3r: final validation of the round: full GPU suite (incl. the reference-built 100k tree and the no-alloc test), smoke, bench + reference arm;
# and whether 16-bit keys (what the key-width policy settles on at 100k uniform / city colliders) cost the traversal anything
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
for kb in 16 24 32; do B2COL_KEY_BITS=$kb timeout 300 python tools/trace_timing.py --workload config2 --colliders 100000 | tail -1; done
for kb in 16 24 32; do B2COL_KEY_BITS=$kb timeout 300 python tools/trace_timing.py --workload config3 --colliders 100000 | tail -1; done
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r3b_bench.json 2> gpurun_out/r3b_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r3b_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r3b_reference.json 2> gpurun_out/r3b_reference.err; echo "reference rc=$?"
python - <<P
import json
d=json.load(open("gpurun_out/r3b_bench.json"))
print("value",d["value"],"e2e",d["e2e"]["value"],"compact",d["e2e"]["compact_results"]["value"], "launches", d.get("gpu_launches"), "traffic", d["roofline"].get("traffic"))
print("rebuild",{k:v for k,v in d["rebuild"].items() if k!="what"})
print("config4", d["config4"]["rebuild"], d["config4"]["refit_only"]["update_ms_median"]); print("config5 update", d["config5"]["update_ms"], d["config5"]["update_frac"])
r=json.load(open("gpurun_out/r3b_reference.json")); print("reference", r["value"])
P
