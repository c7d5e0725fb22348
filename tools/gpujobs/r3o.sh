#!/bin/bash
# r3o: one fence (ticket thread) instead of one per thread at the end of k_refit_keys*
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "adaptive or plan_follows or small_random or config4 or refit or golden" 2>&1 | tail -3
for i in 1 2 3; do timeout 300 python tools/time_update.py; done
