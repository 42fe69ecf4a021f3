#!/bin/bash
# gpurun --timeout 900 -- 'bash tools/gpu_ncu_launches.sh'   : plain bench, then the launch list of the same command under ncu
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_final.log 2>&1 || { tail -5 gpurun_out/bench_final.log; exit 1; }
tail -1 gpurun_out/bench_final.log | cut -c1-400
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log | cut -c1-200
