#!/bin/bash
# sweep of one tuning hook: gpu_sweep.sh VAR "v1 v2 ..." -- run_assign args
VAR=$1; VALS=$2; shift 3
for v in $VALS; do
  echo "$VAR=$v: $(env $VAR=$v timeout 120 python scripts/run_assign.py "$@" 2>&1 | tail -1 | cut -c1-90)"
done
