#!/bin/bash
# parity tests, then the c2 pipeline stage times and its ncu launch list (kernel durations)
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python tools/pipeline_c2.py 8 --json gpurun_out/pipeline_c2.json 2>&1 | tee gpurun_out/pipeline_c2.log | grep -v "B=1 "
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c2.csv python tools/pipeline_c2.py 8 > /dev/null 2>&1
