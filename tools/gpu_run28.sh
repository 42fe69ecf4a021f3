#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_bf16.py -x -q 2>&1 | tail -1
for rep in 1 2; do
AT_B200_LIB=tools/ab/libat_b200_prev.so timeout 120 python tools/tower_time.py 4096 5 2>&1 | tail -1 | sed "s/^/old /"
timeout 120 python tools/tower_time.py 4096 5 2>&1 | tail -1 | sed "s/^/new /"
done | tee gpurun_out/tower_epi.log
