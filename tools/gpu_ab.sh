#!/bin/bash
# gpurun --timeout 900 -- 'bash tools/gpu_ab.sh <variant>'  : A/B of the fused kernel against a variant build + parity + race hunt + timeline
mkdir -p gpurun_out
AT_PREC=f16 timeout 200 python tools/variants.py base "$@" base "$@" | grep tf
timeout 300 python -m pytest tests/test_gpu_bf16.py -m gpu -x -q -k "tf or encoder" 2>&1 | tail -3
timeout 200 python tools/race_hunt.py 300 4133 0 default_tf | tail -3
AT_B200_LIB=$PWD/alphatriangle_b200/libat_b200_exp.so timeout 100 python tools/tf_timeline.py 4096 > gpurun_out/tf_timeline_new.txt 2>&1
tail -4 gpurun_out/tf_timeline_new.txt
