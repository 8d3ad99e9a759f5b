#!/bin/bash
TAG=$1; ENVS=$2; N=$3; D=$4; K=$5
mkdir -p gpurun_out
CMD="env $ENVS python scripts/run_assign.py $N $D $K 3 predict"
$CMD > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:assign_ -s 1 -c 1 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_$TAG.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/plain_$TAG.log | cut -c1-100
