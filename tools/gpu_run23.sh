#!/bin/bash
mkdir -p gpurun_out
show() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:d[k] for k in ('value','ms_per_step','p50_ms','gpu_launches','dtype')}); print(d['e2e'])"; }
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_bf16.log 2>&1; show gpurun_out/bench_bf16.log
AT_NO_ZERO_COPY=1 timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_bf16_copy.log 2>&1; show gpurun_out/bench_bf16_copy.log
timeout 300 python bench.py --workload default_tf --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_tf_4096.log 2>&1; show gpurun_out/bench_tf_4096.log
