#!/bin/bash
# Round bench + the ncu evidence for it.  Outputs under gpurun_out/.
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
cat gpurun_out/bench_full.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench_full.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
# launch list of the same command (short form: 1 timed step, no e2e / cpu legs)
CMD="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/bench_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
# full capture of the dominant kernel at the bench shape
CMD2="python scripts/run_assign.py 10000000 64 100 3"
$CMD2 > gpurun_out/plain_c2full.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:assign_ -s 1 -c 1 -f -o gpurun_out/prof_c2full $CMD2 > gpurun_out/ncu_c2full.log 2>&1
echo "ncu full rc=$?"; cat gpurun_out/plain_c2full.log | cut -c1-120
