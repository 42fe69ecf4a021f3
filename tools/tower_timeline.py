"""Timeline of one board pair inside tower_kernel (experiments build: make -C alphatriangle_b200/csrc experiments).

    AT_B200_LIB=$PWD/alphatriangle_b200/libat_b200_exp.so python tools/tower_timeline.py [batch]

Prints, per warp of CTA 0, the clock64 stamps of the fourth pair that CTA processes, in SM cycles since the earliest stamp.
Tags: issuers 10+l input image ready, 30+l layer l's last slab issued / accumulator committed; epilogue warps 1 pair start,
2 layer 0 written, 50+l accumulator seen complete, 70+l drained and next image posted; producer 100+l first slab of layer l.
"""
import ctypes as C
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
env, mc = at.EnvConfig(), at.ModelConfig(USE_TRANSFORMER=False)
nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=os.environ.get("AT_PREC", "f16"), max_batch=B)
nn.set_weights(synth_state_dict(nn.get_weights(), 1))
base = pack_states(synth_states(256, env, seed=5), env)
packed = torch.from_numpy(np.tile(base, (-(-B // 256), 1))[:B]).cuda()
for _ in range(3):
    nn.engine.evaluate_packed(packed, nn.precision)
torch.cuda.synchronize()
lib = _lib.load()
EV = 256
buf = (C.c_uint64 * (32 * EV))()
lib.at_debug_read_timeline.restype = C.c_int
lib.at_debug_read_timeline.argtypes = [C.c_void_p, C.c_int]
n = lib.at_debug_read_timeline(buf, 32 * EV)
a = np.frombuffer(buf, dtype=np.uint64).reshape(32, EV)
tags, clk = (a >> np.uint64(48)).astype(np.int64), (a & np.uint64(0xFFFFFFFFFFFF)).astype(np.int64)
t0 = clk[tags > 0].min()
names = {0: "producer", 1: "issuer A", 2: "epi A q2 h0", 6: "epi A q2 h1", 10: "epi B q2 h0", 18: "issuer B"}
for w in (0, 1, 18, 2, 6, 10):
    ev = [(int(tags[w, i]), int(clk[w, i] - t0)) for i in range(EV) if tags[w, i] > 0]
    print(f"{names[w]:12s} " + " ".join(f"{t}:{c}" for t, c in ev))
