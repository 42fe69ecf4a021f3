#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py -x -q > gpurun_out/pytest_bf16.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_bf16.log
tail -30 gpurun_out/pytest_bf16.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_bf16.log 2>&1; echo "bench exit $?" >> gpurun_out/bench_bf16.log
tail -3 gpurun_out/bench_bf16.log
