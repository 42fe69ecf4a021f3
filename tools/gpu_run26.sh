#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_bf16.py -x -q 2>&1 | tail -2
for d in 0 4 20 22 16; do
AT_TOWER_DEBUG=$d timeout 120 python tools/tower_time.py 4096 5 2>&1 | tail -1
done | tee gpurun_out/tower_lean.log
