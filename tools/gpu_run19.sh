#!/bin/bash
mkdir -p gpurun_out
python /tmp/tf_time.py > gpurun_out/plain_tf.log 2>&1 || true
cat > /tmp/tf_time.py <<'PY'
import sys; sys.path.insert(0,'.')
import numpy as np, torch
import alphatriangle_b200 as at
from alphatriangle_b200 import _lib
from alphatriangle_b200.packing import pack_states
from alphatriangle_b200.synth import synth_state_dict, synth_states
env, mc = at.EnvConfig(), at.ModelConfig()
nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision="bf16", max_batch=4096)
nn.set_weights(synth_state_dict(nn.get_weights(), 1))
base = pack_states(synth_states(256, env, seed=5), env)
packed = torch.from_numpy(np.tile(base, (16, 1))).cuda()
for _ in range(2):
    nn.engine.evaluate_packed(packed, _lib.PREC_BF16); print(nn.engine.kernel_times())
PY
python /tmp/tf_time.py > gpurun_out/plain_tf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"tower_tf_kernel" -s 1 -c 1 -o gpurun_out/tf_prof2 python /tmp/tf_time.py > gpurun_out/ncu_tf.log 2>&1
tail -2 gpurun_out/ncu_tf.log; cat gpurun_out/plain_tf.log | tail -2
