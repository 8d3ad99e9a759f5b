#!/bin/bash
# ncu captures of the assign kernel (run only after the plain command exits 0).
# usage: gpu_profile.sh TAG N D K
TAG=$1; N=$2; D=$3; K=$4
mkdir -p gpurun_out
CMD="python scripts/run_assign.py $N $D $K 3"
$CMD > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:assign_ -s 1 -c 1 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_$TAG.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/plain_$TAG.log; tail -3 gpurun_out/ncu_$TAG.log
