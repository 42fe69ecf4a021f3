#!/bin/bash
mkdir -p gpurun_out
show() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:d[k] for k in ('value','ms_per_step','p50_ms','gpu_launches','dtype')}); print(d['e2e']); r=d['roofline']; print({k:r[k] for k in ('achieved','frac','kernel_ms','fc_kernel_ms','kernel_share_of_step')})"; }
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 200 python tools/feature_bw.py 2>&1 | tail -4 | tee gpurun_out/feature_bw.log
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_bf16.log 2>&1; show gpurun_out/bench_bf16.log
timeout 300 python bench.py --workload default_tf --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_tf_4096.log 2>&1; show gpurun_out/bench_tf_4096.log
