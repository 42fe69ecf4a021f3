#!/bin/bash
mkdir -p gpurun_out
for mode in 0 3; do for d in 0 4; do AT_TOWER_MODE=$mode AT_TOWER_DEBUG=$d timeout 120 python tools/tower_time.py 4096 3 2>&1 | tail -1 | sed "s/^/mode=$mode /"; done; done | tee gpurun_out/tower_modes3.log
AT_TOWER_MODE=3 timeout 300 python -m pytest tests/test_gpu_bf16.py -x -q -k conv_only 2>&1 | tail -3
