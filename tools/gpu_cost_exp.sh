#!/bin/bash
# run every tools/bin/exp_cost_* harness binary; ARGS = harness arguments
mkdir -p gpurun_out
for b in tools/bin/exp_cost_*; do
  for a in "${@:-32}"; do echo "== $b $a"; timeout 300 $b $a | tail -2; done
done 2>&1 | tee gpurun_out/exp_cost.log
