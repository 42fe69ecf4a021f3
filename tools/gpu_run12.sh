#!/bin/bash
mkdir -p gpurun_out
show() { tail -1 $1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:d[k] for k in ('value','ms_per_step','p50_ms','gpu_launches','dtype')}); print(d['e2e']); r=d['roofline']; print({k:r[k] for k in ('achieved','frac','kernel','kernel_ms','fc_kernel_ms','kernel_share_of_step')})"; }
timeout 300 python bench.py --workload default_tf --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_tf_4096.log 2>&1; show gpurun_out/bench_tf_4096.log
timeout 300 python bench.py --workload default_tf --batch 16384 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tf_16384.log 2>&1; show gpurun_out/bench_tf_16384.log
timeout 300 python bench.py --workload default_tf --precision fp32 --batch 4096 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tf_fp32.log 2>&1; show gpurun_out/bench_tf_fp32.log
timeout 300 python bench.py --workload conv_only --batch 16384 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_conv_16384.log 2>&1; show gpurun_out/bench_conv_16384.log
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
