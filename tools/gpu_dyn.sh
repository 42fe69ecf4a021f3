#!/bin/bash
# gpurun --timeout 900 -- 'bash tools/gpu_dyn.sh <variant>'  : both tower kernels against a variant build + the tests that exercise them + race hunt
mkdir -p gpurun_out
AT_PREC=f16 timeout 300 python tools/variants.py base "$@" base "$@"
timeout 500 python -m pytest tests/test_gpu_bf16.py tests/test_zz_no_alloc.py tests/test_gpu_publish.py -m gpu -x -q 2>&1 | tail -2
for B in 4133 297 31; do timeout 200 python tools/race_hunt.py 300 $B 0 default_tf,conv_only | tail -4; done
