#!/bin/bash
# run every tools/exp_sweep_* harness binary (real geometry: 72 bands of 1080 rows); CASES = "B:publish-interval" pairs
mkdir -p gpurun_out
CASES=${CASES:-"8:16 32:16"}
for b in tools/exp_sweep_*; do
  for c in $CASES; do
    B=${c%%:*}; S=${c##*:}
    echo "== $b B=$B pub=$S"; timeout 120 $b $B 1080 0 0 $S | grep pass
  done
done 2>&1 | tee gpurun_out/exp.log
