#!/bin/bash
# One gpurun call: GPU parity tests + bench (both arms) + stage times.  No profiler.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; tail -1 gpurun_out/bench.log
python tools/stage_times.py > gpurun_out/stage_times.log 2>&1; cat gpurun_out/stage_times.log
