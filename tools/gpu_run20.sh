#!/bin/bash
mkdir -p gpurun_out
for d in 0 16 32 48 64 128 192; do AT_TOWER_DEBUG=$d timeout 120 python tools/tf_time.py 4096 2>&1 | tail -1; done | tee gpurun_out/tf_variants.log
