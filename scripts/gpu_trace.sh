#!/bin/bash
# timeline of CTA 0 (trace build in scripts/probes/lib_trace.so): gpu_trace.sh "ENV=.. ENV=.." N D K [predict]
mkdir -p gpurun_out
ENVS=$1; shift
W=$(env $ENVS python scripts/run_assign.py "$@" | tee /dev/stderr | sed 's/.*mma_warps=\([0-9]*\).*/\1/')
env $ENVS FMLK_LIB_PATH=$PWD/scripts/probes/lib_trace.so FMLK_TRACE_FILE=gpurun_out/trace.bin python scripts/run_assign.py "$@" | cut -c1-100
python scripts/trace_timeline.py gpurun_out/trace.bin $W | awk 'NR<=2 || /acc\)/ && NR<16 || /scheduler 0/'
