#!/bin/bash
# plain run first, then one ncu --set full capture of the second evaluate step (tower, fc, heads)
mkdir -p gpurun_out
python tools/tower_time.py 4096 2 > gpurun_out/plain_step.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"tower_kernel|fc_kernel|heads_mma" -s 3 -c 3 -o gpurun_out/step_prof4 python tools/tower_time.py 4096 2 > gpurun_out/ncu_step4.log 2>&1
tail -2 gpurun_out/ncu_step4.log; tail -2 gpurun_out/plain_step.log
