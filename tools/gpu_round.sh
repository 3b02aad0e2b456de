#!/bin/bash
# One gpurun call: GPU parity tests, bench, per-stage times, ncu launch list and one full capture of the sweep.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>&1; tail -1 gpurun_out/bench.log
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1; tail -1 gpurun_out/bench_ref.log
python tools/stage_times.py > gpurun_out/stage_times.log 2>&1; cat gpurun_out/stage_times.log
python bench.py --batch 2 --steps 1 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
    python bench.py --batch 2 --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1
python bench.py --batch 2 --steps 1 --warmup 3 --no-cpu > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_sweep -s 8 -c 2 -o gpurun_out/sweep_full \
    python bench.py --batch 2 --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
