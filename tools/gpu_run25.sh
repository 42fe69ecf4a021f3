#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
AT_HEADS_MODE=0 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_heads0.log 2>&1; tail -1 gpurun_out/bench_heads0.log | cut -c1-400
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_heads1.log 2>&1; tail -1 gpurun_out/bench_heads1.log | cut -c1-400
python bench.py --workload default_tf --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_tf_heads1.log 2>&1; tail -1 gpurun_out/bench_tf_heads1.log | cut -c1-400
