#!/bin/bash
# Round-2 ncu evidence.  Usage: gpu_prof_r2.sh <tag> [B]
#  1. launch list of one c3 step at batch B (gpu__time_duration only, no replay)
#  2. --set full capture of k_cost_fused + both k_sweep launches of one step at batch B
#  3. dram__bytes + time of the HBM-bound stages (k_remap_u8, k_planes, k_bgr2gray, k_reproject) in the c2/1080p pipeline
TAG=${1:-x}; B=${2:-32}
mkdir -p gpurun_out
CMD="python bench.py --batch $B --steps 1 --warmup 3 --no-cpu --no-other"
$CMD > gpurun_out/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$TAG.csv $CMD > /dev/null 2>&1
echo "launch list rc=$?"
# per step: k_planes, k_cost_fused, k_sweep x2, ...; warm-up 3 steps + 2 e2e warm-ups + 1 stage-timed step come first
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_sweep|k_cost_fused' -s 15 -c 3 -o gpurun_out/full_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full capture rc=$?"; tail -2 gpurun_out/ncu_full_$TAG.log
python tools/pipeline_c2.py 8 > gpurun_out/pipeline_plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:'k_remap|k_planes|k_bgr2gray|k_reproject|k_init_undistort' --csv --log-file gpurun_out/hbm_stages_$TAG.csv \
    python tools/pipeline_c2.py 8 > /dev/null 2>&1
echo "hbm stages rc=$?"
ls -la gpurun_out | tail -8
