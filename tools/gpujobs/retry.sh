#!/bin/bash
# dev: keep asking for the GPU until the job ran: tools/gpujobs/retry.sh <log> <gpurun args...>
log=$1; shift
for i in $(seq 1 40); do
  timeout 3400 gpurun "$@" > "$log" 2>&1
  if ! grep -q "status=transient" "$log"; then exit 0; fi
  sleep 20
done
