#!/bin/bash
# A/B: scripts/probes/lib_base.so (previous build) against the in-tree build, then parity tests.
mkdir -p gpurun_out
SHAPES=("10000000 64 100 5" "10000000 64 1000 3" "4000000 128 100 3" "4000000 32 100 3" "2000000 200 50 3")
for a in "${SHAPES[@]}"; do
  echo "base: $(FMLK_LIB_PATH=$PWD/scripts/probes/lib_base.so timeout 120 python scripts/run_assign.py $a 2>&1 | tail -1 | cut -c1-200)"
  echo "new : $(timeout 120 python scripts/run_assign.py $a 2>&1 | tail -1 | cut -c1-200)"
done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
