#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for m in 0 1 2; do
AT_HEADS_MODE=$m python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_heads$m.log 2>&1; tail -1 gpurun_out/bench_heads$m.log | cut -c1-330
done
python bench.py --workload default_tf --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_tf_heads2.log 2>&1; tail -1 gpurun_out/bench_tf_heads2.log | cut -c1-330
