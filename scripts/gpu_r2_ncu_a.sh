#!/bin/bash
# launch list of the short bench (kernel shares of the step), one ncu command per call
mkdir -p gpurun_out
SMALL="bench.py --steps 1 --warmup 1 --prompts 296 --batch-prompts 296 --no-cpu-baseline --no-workloads"
python $SMALL > gpurun_out/plain_small.log 2>&1 && echo "plain ok" &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02zv_launches_bench_296prompts.csv python $SMALL > gpurun_out/ncu_launch.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/r02zv_launches_bench_296prompts.csv
