#!/bin/bash
# ncu: launch list (one step) + full captures of the cost / sweep / ccl kernels.  Usage: gpu_prof.sh <tag> <batch>
TAG=${1:-x}; B=${2:-4}
mkdir -p gpurun_out
CMD="python bench.py --batch $B --steps 1 --warmup 3 --no-cpu"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$TAG.csv $CMD > /dev/null 2>&1
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_sweep|k_cost' -s 9 -c 3 -o gpurun_out/full_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
tail -2 gpurun_out/ncu_full_$TAG.log; tail -1 gpurun_out/plain_$TAG.log | cut -c1-300
