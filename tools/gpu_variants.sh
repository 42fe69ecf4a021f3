#!/bin/bash
# gpurun --timeout 900 -- 'bash tools/gpu_variants.sh name1 name2 ...'   (A/B timing of prebuilt libat_b200_<name>.so variants)
mkdir -p gpurun_out
python tools/variants.py "$@" 2>&1 | tee gpurun_out/variants_$(date +%H%M%S).log
