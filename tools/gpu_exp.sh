#!/bin/bash
# run every tools/exp_sweep_* harness binary at B=8 and B=32 (real geometry: 72 bands of 1080 rows)
mkdir -p gpurun_out
for b in tools/exp_sweep_*; do
  for B in 8 32; do
    echo "== $b B=$B"; timeout 120 $b $B 1080 16 | grep pass
  done
done 2>&1 | tee gpurun_out/exp.log
