#!/bin/bash
mkdir -p gpurun_out
AT_TOWER_PAIR=0 timeout 120 python tools/tower_time.py 4096 4 2>&1 | tail -1
AT_TOWER_PAIR=1 timeout 120 python tools/tower_time.py 4096 4 2>&1 | tail -3
timeout 600 python -m pytest tests/test_gpu_bf16.py -x -q 2>&1 | tail -8
