#!/bin/bash
# plain run first, then one ncu --set full capture of the second evaluate step (features, tower, fc, heads)
mkdir -p gpurun_out
python tools/tower_time.py 4096 2 > gpurun_out/plain_step.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"features|tower_kernel|fc_kernel|heads_fused" -s 4 -c 4 -o gpurun_out/step_prof3 python tools/tower_time.py 4096 2 > gpurun_out/ncu_step3.log 2>&1
tail -2 gpurun_out/ncu_step3.log; tail -2 gpurun_out/plain_step.log
