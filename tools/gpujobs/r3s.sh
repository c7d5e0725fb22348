#!/bin/bash
# r3s: compute-sanitizer over the steady-state frame's kernels (new this round: bulk-copy refit, climber compaction, static scratch)
timeout 120 python tools/sanitize_small.py
for tool in memcheck racecheck synccheck; do
  echo "== $tool"
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_small.py 2>&1 | grep -v "^$" | tail -12
done
