#!/bin/bash
# c2 pipeline: stage times + HBM rooflines, ncu launch list, full captures of the rectify / reproject kernels
mkdir -p gpurun_out
python tools/pipeline_c2.py 8 --json gpurun_out/pipeline_c2.json 2>&1 | tee gpurun_out/pipeline_c2.log
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c2.csv python tools/pipeline_c2.py 8 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_remap_u8|k_reproject|k_init_undistort|k_bgr2gray|k_box_median' -c 12 -o gpurun_out/full_c2 python tools/pipeline_c2.py 8 > gpurun_out/ncu_full_c2.log 2>&1
tail -2 gpurun_out/ncu_full_c2.log
