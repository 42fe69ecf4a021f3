#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
timeout 600 python tools/selfplay_pattern.py 128 40 32 conv_only 2>&1 | tail -2 | tee gpurun_out/selfplay_pattern.log
timeout 600 python tools/selfplay_pattern.py 128 40 32 default_tf 2>&1 | tail -2 | tee -a gpurun_out/selfplay_pattern.log
