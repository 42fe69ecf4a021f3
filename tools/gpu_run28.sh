#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py -x -q 2>&1 | tail -1
for rep in 1 2; do
AT_B200_LIB=tools/ab/libat_b200_mode3.so timeout 120 python tools/tower_time.py 4096 5 2>&1 | tail -1 | sed "s/^/old /"
for d in 0 64; do
AT_TOWER_DEBUG=$d timeout 120 python tools/tower_time.py 4096 5 2>&1 | tail -1 | sed "s/^/new /"
done; done | tee gpurun_out/tower_ab2.log
