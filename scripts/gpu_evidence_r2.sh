#!/bin/bash
# Round-2 evidence in one gpurun call: the bench lines, the launch list of the bench, and one
# `ncu --set full` capture per kernel family at the shapes DESIGN.md quotes.  Everything lands under
# gpurun_out/; scripts/refresh_profiles_r2.py turns it into the tracked summaries under profiles/.
# usage: gpu_evidence_r2.sh 1|2|3   (separate calls: gpurun merges at most 64 MiB of gpurun_out/ per call)
PART=${1:-1}
mkdir -p gpurun_out
if [ "$PART" = 1 ]; then
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference_n1.json 2>> gpurun_out/r2_bench_n1.err; echo "ref rc=$?"
python bench.py --fast --steps 5 --warmup 3 --no-north-star --no-extra --no-cpu > gpurun_out/r2_bench_fast_n1.json 2>> gpurun_out/r2_bench_n1.err; echo "fast rc=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-north-star --no-extra"
$CMD > gpurun_out/bench_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_bench.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
fi
cap() { # name, kernel regex, env, args...
  local name=$1 regex=$2 envs=$3; shift 3
  env $envs python scripts/run_assign.py "$@" > gpurun_out/plain_$name.log 2>&1 && \
  env $envs ncu --set full --clock-control none --import-source on -k regex:$regex -s 1 -c 1 -f -o gpurun_out/r2_$name python scripts/run_assign.py "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$? $(tail -1 gpurun_out/plain_$name.log | cut -c1-80)"
}
if [ "$PART" = 3 ]; then
cap c1shape_train_small assign_small FMLK_FAST=0 8000000 10 2 3
cap c2_train_ordered assign_kernel FMLK_DET=1 10000000 64 100 3
elif [ "$PART" = 1 ]; then
cap c2_train_f64 assign_kernel FMLK_FAST=0 10000000 64 100 3
cap c2_train_fast assign_fast FMLK_FAST=1 10000000 64 100 3
cap c2_predict_fast assign_fast FMLK_FAST=1 10000000 64 100 3 predict
cap c1b_train_f64 assign_kernel FMLK_FAST=0 1000000 100 10 4
cap c1b_predict_f64 assign_kernel FMLK_FAST=0 1000000 100 10 4 predict
else
cap c4shape_train_f64 assign_ring FMLK_FAST=0 10000000 64 1000 2
cap c4shape_train_fast assign_fast FMLK_FAST=1 10000000 64 1000 2
cap c3shape_predict_f64 assign_ring FMLK_FAST=0 4000000 128 1000 2 predict
cap c5shape_train_f64 assign_ring FMLK_FAST=0 4000000 128 256 2
# KnnModel.transform (scripts/run_knn.py: 65536 rows x 64 against 10000 training points, k = 5)
python scripts/run_knn.py > gpurun_out/plain_knn.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:knn_predict -s 1 -c 1 -f -o gpurun_out/r2_knn python scripts/run_knn.py > gpurun_out/ncu_knn.log 2>&1
echo "knn rc=$?"
python scripts/measure_knn.py > gpurun_out/knn_measure.log 2>&1; echo "knn measure rc=$?"
python scripts/measure_configs.py c1 c1b c2 c3 c4 c5 --no-cpu > gpurun_out/r2_configs_n1.jsonl 2> gpurun_out/r2_configs.err; echo "configs rc=$?"
./scripts/probes/l2_red_bench > gpurun_out/r2_l2_add_bench.txt 2>&1; echo "l2 bench rc=$?"
fi
