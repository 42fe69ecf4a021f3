#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
tail -2 gpurun_out/ncu_bench.log | cut -c1-300
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/launches.csv')) if len(r)>10]
hdr=rows[0]; i_name=hdr.index('Kernel Name'); i_val=hdr.index('Metric Value'); i_unit=hdr.index('Metric Unit')
agg={}
for r in rows[1:]:
    n=r[i_name][:70]; v=float(r[i_val].replace(',','')); u=r[i_unit]
    a=agg.setdefault(n,[0,0.0,u]); a[0]+=1; a[1]+=v
for n,(c,t,u) in sorted(agg.items(), key=lambda kv:-kv[1][1]): print(f"{c:4d} x {t/c:12.1f} {u}  {n}")
PY
