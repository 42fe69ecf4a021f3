#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py -x -q 2>&1 | tail -3
for d in 0 1 2 4 5 6 8; do AT_TOWER_DEBUG=$d timeout 120 python tools/tower_time.py 4096 4 2>&1 | tail -1; done | tee gpurun_out/tower_variants2.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_bf16.log 2>&1; tail -1 gpurun_out/bench_bf16.log | cut -c1-400
