#!/bin/bash
# gpurun --timeout 600 -- 'bash tools/gpu_ncu_step.sh [conv_only|default_tf]'
# plain run first, then ONE ncu --set full capture of the second evaluate step (tower, fc, heads); read the report here with
#   ncu -i gpurun_out/step_prof.ncu-rep --page raw --csv      /      --page source --csv
mkdir -p gpurun_out
W=${1:-conv_only}
if [ "$W" = default_tf ]; then CMD="python tools/tf_time.py 4096"; K='tower_tf_kernel|fc_kernel|heads_mma'; else CMD="python tools/tower_time.py 4096 2"; K='tower_kernel|fc_kernel|heads_mma'; fi
$CMD > gpurun_out/plain_step.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"$K" -s 3 -c 3 -o gpurun_out/step_prof $CMD > gpurun_out/ncu_step.log 2>&1
tail -2 gpurun_out/ncu_step.log; tail -2 gpurun_out/plain_step.log
