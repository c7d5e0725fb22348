#!/bin/bash
# r2m: key statistic counters out of the header (k_build regression fix), run-length statistic; aimed rays vs key width on clustered scenes
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or build or refit or golden or config4 or degenerate" 2>&1 | tail -4
echo "== default (adaptive keys)"; python tools/time_update.py
echo "== B2COL_ADAPTIVE_KEYS=0"; B2COL_ADAPTIVE_KEYS=0 python tools/time_update.py
echo "== phases (static width)"; B2COL_ADAPTIVE_KEYS=0 B2COL_SORT=lsd python tools/phase_timing.py
for kb in 32 40 48; do echo "== config5 2M aimed KEY_BITS=$kb"; B2COL_KEY_BITS=$kb python tools/trace_timing.py --workload config5 --colliders 2000000 --aimed | tail -1; done
for kb in 40 48 56; do echo "== config5 10M aimed KEY_BITS=$kb"; B2COL_KEY_BITS=$kb python tools/trace_timing.py --workload config5 --colliders 10000000 --aimed | tail -1; done
for kb in 24 32; do echo "== config3 aimed KEY_BITS=$kb"; B2COL_KEY_BITS=$kb python tools/trace_timing.py --workload config3 --aimed | tail -1; done
