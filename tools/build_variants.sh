#!/bin/bash
# dev tool: build kernel variants of libb2col.so into variants/ (git-ignored): name=defines ...
#   tools/build_variants.sh "mb8=-DB2_TRACE_MIN_BLOCKS=8" "st1=-DB2_TRACE_STEPS=1"
cd "$(dirname "$0")/.."
mkdir -p variants
for spec in "$@"; do
  name="${spec%%=*}"; defs="${spec#*=}"
  B2COL_DEFINES="$defs" python -m engine_b200.build --out=variants/libb2col_$name.so > /dev/null &
done
wait
ls -la variants/
