#!/bin/bash
mkdir -p gpurun_out
for d in 0 1 2 3 4 6 7 8 15; do AT_TOWER_DEBUG=$d timeout 120 python tools/tower_time.py 4096 4 2>&1 | tail -1; done | tee gpurun_out/tower_variants.log
python tools/tower_time.py 4096 2 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tower_kernel -c 1 -o gpurun_out/tower_prof python tools/tower_time.py 4096 2 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
