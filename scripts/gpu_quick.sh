#!/bin/bash
# Quick pass: parity tests, the kernel at the main shapes, the e2e breakdown.  Outputs under gpurun_out/.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
for a in "10000000 64 100 5" "10000000 64 100 5 predict" "10000000 64 1000 3" "1000000 100 10 10" "1000000 100 10 10 predict" "2000000 10 2 10"; do
  timeout 120 python scripts/run_assign.py $a 2>&1 | tail -1 | cut -c1-250
done
timeout 300 python scripts/e2e_breakdown.py 2>&1 | tail -40
