#!/bin/bash
mkdir -p gpurun_out
python tools/tower_time.py 4096 2 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"heads_fused|tower_kernel|fc_kernel|features_kernel" -s 4 -c 4 -o gpurun_out/step_prof python tools/tower_time.py 4096 2 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
