#!/bin/bash
# compute-sanitizer memcheck + racecheck + synccheck over tools/sanitize_run.py; logs -> gpurun_out/
mkdir -p gpurun_out
python tools/sanitize_run.py > gpurun_out/sanitize_plain.log 2>&1; echo "plain rc=$?"
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --log-file gpurun_out/sanitize_$tool.log python tools/sanitize_run.py > gpurun_out/sanitize_${tool}_stdout.log 2>&1
  echo "$tool rc=$?"; tail -3 gpurun_out/sanitize_$tool.log
done
