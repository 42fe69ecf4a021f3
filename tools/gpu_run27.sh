#!/bin/bash
mkdir -p gpurun_out
python tools/tower_time.py 4096 2 > gpurun_out/plain_step.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"tower_kernel" -s 1 -c 1 -o gpurun_out/tower_lean_prof python tools/tower_time.py 4096 2 > gpurun_out/ncu_tower_lean.log 2>&1
tail -2 gpurun_out/ncu_tower_lean.log; tail -1 gpurun_out/plain_step.log
