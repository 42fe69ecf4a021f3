#!/bin/bash
# gpurun --timeout 900 -- 'bash tools/gpu_race.sh'    (run-to-run determinism of every tensor-core path, many launches)
mkdir -p gpurun_out
{
  python tools/race_hunt.py 600 4133 0 default_tf_sharp,conv_only_peaked,default_tf,conv_only
  python tools/race_hunt.py 200 16384 0 default_tf,conv_only
  python tools/race_hunt.py 1000 296 0 default_tf,conv_only
  python tools/race_hunt.py 1000 31 0 default_tf,conv_only
  python tools/race_hunt.py 400 4096 0 tf_1layer_ff128_gelu,tf_3layers_ff512,tf_cuda_core_heads,conv_unfused_tail
} 2>&1 | grep -v "^$" | tee gpurun_out/race_hunt.log | tail -60
