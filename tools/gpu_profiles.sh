#!/bin/bash
# gpurun --timeout 1800 -- 'bash tools/gpu_profiles.sh'
# Evidence for profiles/: latency sweep, launch list of the bench command, one ncu --set full capture per BASELINE model
# (raw page exported here as CSV + the kernel-source hash it belongs to, which bench.py checks before quoting its DRAM traffic).
mkdir -p gpurun_out
python tools/latency_sweep.py > gpurun_out/latency_sweep.jsonl 2> gpurun_out/latency_sweep.err || tail -5 gpurun_out/latency_sweep.err
grep '"batch": 32,' gpurun_out/latency_sweep.jsonl | cut -c1-260
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --no-api --sustained 0 > gpurun_out/plain_bench_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-workloads --no-api --sustained 0 > gpurun_out/ncu_launches.log 2>&1
for W in conv_only default_tf; do
  if [ "$W" = default_tf ]; then CMD="python tools/tf_time.py 4096"; K='tower_tf_kernel|fc_kernel|heads_mma'; else CMD="python tools/tower_time.py 4096 2"; K='tower_kernel|fc_kernel|heads_mma'; fi
  AT_PREC=auto $CMD > gpurun_out/plain_$W.log 2>&1 && \
  AT_PREC=auto ncu --set full --clock-control none --import-source on -k regex:"$K" -s 3 -c 3 -o gpurun_out/prof_$W -f $CMD > gpurun_out/ncu_$W.log 2>&1
  tail -1 gpurun_out/plain_$W.log
  ncu -i gpurun_out/prof_$W.ncu-rep --page raw --csv > gpurun_out/r02_ncu_full_${W}_b4096_raw.csv 2>/dev/null
  python -c "import bench, json; print(json.dumps({'kernel_source_hash': bench.kernel_source_hash(), 'command': '$CMD', 'batch': 4096}))" > gpurun_out/r02_ncu_full_${W}_b4096_raw.json
done
