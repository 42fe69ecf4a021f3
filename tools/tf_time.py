"""Time the fused tower+transformer kernel (AT_TOWER_DEBUG variants: 16 no encoder MMAs, 32 no encoder epilogue math, 64 no softmax,
128 no LayerNorm)."""
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import alphatriangle_b200 as at  # noqa: E402
from alphatriangle_b200 import _lib  # noqa: E402
from alphatriangle_b200.packing import pack_states  # noqa: E402
from alphatriangle_b200.synth import synth_state_dict, synth_states  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
env, mc = at.EnvConfig(), at.ModelConfig()
nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=os.environ.get("AT_PREC", "bf16"), max_batch=B)
nn.set_weights(synth_state_dict(nn.get_weights(), 1))
base = pack_states(synth_states(256, env, seed=5), env)
packed = torch.from_numpy(np.tile(base, (-(-B // 256), 1))[:B]).cuda()
nn.engine.set_kernel_timing(True)
ts = []
for _ in range(6):
    probs, value, _flags = nn.engine.evaluate_packed(packed, nn.precision)
    ts.append(nn.engine.kernel_times()["tower_ms"])
chk = f"{float(value.double().sum()):.9f}/{float((probs.double() * torch.arange(probs.shape[1], device=probs.device)).sum()):.6f}"
print(f"{_lib.PRECISION_NAMES[nn.precision]} debug={os.environ.get('AT_TOWER_DEBUG', '0')} B={B} tower_tf_ms={min(ts):.4f} chk={chk}")
