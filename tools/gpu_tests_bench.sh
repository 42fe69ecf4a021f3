#!/bin/bash
# gpurun --timeout 1500 -- 'bash tools/gpu_tests_bench.sh'   : GPU parity tests, then both bench workloads (plain runs, no profiler)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_conv_only.json 2> gpurun_out/bench_conv_only.err || tail -5 gpurun_out/bench_conv_only.err
python bench.py --workload default_tf --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_default_tf.json 2> gpurun_out/bench_default_tf.err || tail -5 gpurun_out/bench_default_tf.err
for f in gpurun_out/bench_conv_only.json gpurun_out/bench_default_tf.json; do
  tail -1 $f | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['config']['workload'][:40], round(d['value']), 'e2e', round(d['e2e']['value']), 'roofline', round(d['roofline']['frac'],3))"
done
