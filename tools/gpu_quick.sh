#!/bin/bash
# quick GPU check: parity tests + per-stage times
mkdir -p gpurun_out
timeout 120 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 120 python tools/stage_times.py > gpurun_out/stage_times.log 2>&1; cat gpurun_out/stage_times.log
