#!/bin/bash
# Development aid (GPU box): run tools/quick_bench.py under a list of environment settings.
# usage: tools/sweep_env.sh "DIS_EARLY_STEP=0.1" "DIS_EARLY_STEP=0.3 DIS_HANDOFF=24,48,96,192" ...
cd "$(dirname "$0")/.."
for spec in "$@"; do
  echo "== $spec"
  env $spec python tools/quick_bench.py 3 2>&1 | tail -4
done
