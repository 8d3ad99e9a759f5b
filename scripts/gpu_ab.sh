#!/bin/bash
# A/B: every scripts/probes/lib_*.so against the in-tree build (train + predict shapes), then parity tests.
mkdir -p gpurun_out
SHAPES=("10000000 64 100 5" "10000000 64 100 5 predict" "10000000 64 1000 3" "4000000 32 100 3" "2000000 200 50 3")
for a in "${SHAPES[@]}"; do
  for l in scripts/probes/lib_*.so; do
    echo "$(basename $l .so): $(FMLK_LIB_PATH=$PWD/$l timeout 120 python scripts/run_assign.py $a 2>&1 | tail -1 | cut -c1-100)"
  done
  echo "in-tree: $(timeout 120 python scripts/run_assign.py $a 2>&1 | tail -1 | cut -c1-200)"
done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
