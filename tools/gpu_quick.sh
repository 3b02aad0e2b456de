#!/bin/bash
# quick GPU check: short smoke first (a hung kernel costs 40 s, not minutes), then parity tests + per-stage times
mkdir -p gpurun_out
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1 || { echo "SMOKE FAILED rc=$?"; tail -5 gpurun_out/smoke.log; exit 1; }
timeout 150 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 120 python tools/stage_times.py > gpurun_out/stage_times.log 2>&1; cat gpurun_out/stage_times.log
