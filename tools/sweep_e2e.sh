#!/bin/bash
# Development aid (GPU box): host-to-host throughput under a list of DIS_OUT_SLICES settings ("" = built-in default).
cd "$(dirname "$0")/.."
for s in "${@:-default 3,1 4,1 5,2 4 8,4,2,1}"; do
  for t in $s; do
    echo "== DIS_OUT_SLICES=$t"
    if [ "$t" = default ]; then python tools/e2e_bench.py 10 2>&1 | tail -1; else DIS_OUT_SLICES=$t python tools/e2e_bench.py 10 2>&1 | tail -1; fi
  done
done
