#!/bin/bash
# where does the d-blocked kernel beat the main kernel?  usage: gpu_wide_ab.sh "N D K R [predict]" ...
for a in "$@"; do
  for v in 0 1; do
    echo "wide=$v: $(FMLK_TUNE_WIDE=$v timeout 200 python scripts/run_assign.py $a 2>&1 | tail -1 | cut -c1-86)"
  done
done
