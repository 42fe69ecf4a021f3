#!/bin/bash
mkdir -p gpurun_out
for m in 3; do
for d in 0 4 8196 8192 8198; do
AT_TOWER_MODE=$m AT_TOWER_DEBUG=$d timeout 120 python tools/tower_time.py 4096 5 2>&1 | tail -1 | sed "s/^/mode=$m /"
done; done | tee gpurun_out/tower_modes5e.log
