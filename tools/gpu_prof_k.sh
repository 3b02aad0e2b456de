#!/bin/bash
# ncu full capture of selected kernels.  Usage: gpu_prof_k.sh <tag> <batch> <kernel-regex> <skip> <count>
TAG=${1:-x}; B=${2:-4}; RX=${3:-k_cost}; SK=${4:-3}; CN=${5:-1}
mkdir -p gpurun_out
CMD="python bench.py --batch $B --steps 1 --warmup 3 --no-cpu"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$RX" -s $SK -c $CN -o gpurun_out/full_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
tail -2 gpurun_out/ncu_full_$TAG.log; tail -1 gpurun_out/plain_$TAG.log | cut -c1-300
