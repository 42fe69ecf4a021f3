#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 120 python tools/tower_time.py 4096 4 2>&1 | tail -1
timeout 300 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_bf16.log 2>&1; tail -1 gpurun_out/bench_bf16.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print({k:d[k] for k in ('value','ms_per_step','p50_ms','gpu_launches')}); print(d['e2e']); print(d['roofline']); print(d['clocks']); print(d.get('cpu_baseline'))"
