#!/bin/bash
# parity tests + c3 stage times at B=8 and B=32 (no bench, no profiler)
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python tools/stage_times.py 8; python tools/stage_times.py 32
