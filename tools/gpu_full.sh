#!/bin/bash
# gpurun --timeout 2400 -- 'bash tools/gpu_full.sh'   : the whole GPU suite, smoke(), and the default bench line (plain runs)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r2.json 2> gpurun_out/bench_r2.err || tail -20 gpurun_out/bench_r2.err
tail -1 gpurun_out/bench_r2.json | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('value', round(d['value']), d['dtype'], 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'], 3), 'sustained', round(d['sustained']['states_per_s']))
for k, w in d.get('workloads', {}).items(): print(k, round(w['states_per_s']), 'e2e', round(w['e2e']['value']), 'frac', round(w['roofline']['frac'], 3), 'kernel_ms', round(w['roofline']['kernel_ms'], 4))
print('api', round(d['api']['value']), 'cpu', round(d['cpu_baseline']['value']), d['cpu_baseline']['kind'])
"
