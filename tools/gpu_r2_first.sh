#!/bin/bash
# gpurun --timeout 2400 -- 'bash tools/gpu_r2_first.sh'
# round-2 first pass: GPU parity tests (verbose parity prints kept), bench (plain), launch list under ncu, one ncu --set full
# capture per BASELINE model (each only after its plain command exited 0)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -s -rA 2>&1 | tee gpurun_out/pytest_gpu.log | grep -E "bf16 parity|passed|failed|FAILED|ERROR" | tail -40
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err || { tail -20 gpurun_out/bench_r2.err; }
tail -1 gpurun_out/bench_r2.json | cut -c1-600
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --no-api --sustained 0 > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log | cut -c1-200
for W in conv_only default_tf; do
  if [ "$W" = default_tf ]; then CMD="python tools/tf_time.py 4096"; K='tower_tf_kernel|fc_kernel|heads_mma'; else CMD="python tools/tower_time.py 4096 2"; K='tower_kernel|fc_kernel|heads_mma'; fi
  $CMD > gpurun_out/plain_$W.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:"$K" -s 3 -c 3 -o gpurun_out/prof_$W $CMD > gpurun_out/ncu_$W.log 2>&1
  tail -1 gpurun_out/plain_$W.log; tail -1 gpurun_out/ncu_$W.log | cut -c1-200
done
