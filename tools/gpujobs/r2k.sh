#!/bin/bash
# r2k: sanity of the restored tree (full GPU suite) + Morton key width A/B (24 / 27 / 30 / 32 bits): update time vs trace rate
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for kb in 0 24 27 30; do
  echo "== B2COL_KEY_BITS=$kb"
  B2COL_KEY_BITS=$kb python tools/time_update.py
  for w in "config3 1000000" "config5 2000000" "config2 100000" "city 1000000"; do set -- $w; B2COL_KEY_BITS=$kb python tools/trace_timing.py --workload $1 --colliders $2 | tail -1; done
done
echo "== phases default"; python tools/phase_timing.py
echo "== phases 24"; B2COL_KEY_BITS=24 python tools/phase_timing.py
python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2k_bench.json 2> gpurun_out/r2k_bench.err; echo "bench rc=$?"; cat gpurun_out/r2k_bench.json | head -c 1500
