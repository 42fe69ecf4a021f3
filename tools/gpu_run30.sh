#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_pipe.log 2>&1; tail -1 gpurun_out/bench_pipe.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e'])"
python bench.py --workload default_tf --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_tf_pipe.log 2>&1; tail -1 gpurun_out/bench_tf_pipe.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e'])"
