"""What one big evaluation costs when something else holds a few SMs (a stand-in for the NCCL kernel of a weight publication).

    AT_B200_LIB=... python tools/contention.py [batch] [hog] [hog_cycles_per_step]

With hog_cycles_per_step: `hog` CTAs of the probe library's hog kernel (each a whole SM) spin on a side stream for the whole
timed region. Without it:
A second engine evaluates `hog_batch` boards (hog_batch / 2 CTAs of the same tower kernel, i.e. that many SMs) in a loop on a
side stream while the main stream times evaluations of `batch` boards. With board pairs assigned by blockIdx stride the big
launch's CTAs that find no free SM run as a second wave; with pairs handed out by a counter they just take fewer pairs.
"""
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import alphatriangle_b200 as at  # noqa: E402
from alphatriangle_b200 import _lib  # noqa: E402
from alphatriangle_b200.packing import pack_states  # noqa: E402
from alphatriangle_b200.synth import synth_state_dict, synth_states  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
HOG = int(sys.argv[2]) if len(sys.argv) > 2 else 32
HOG_CYCLES = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0  # > 0: the probe library's hog kernel, this many SM cycles per step
prec = os.environ.get("AT_PREC", "f16")
probe = _lib.load_probe() if HOG_CYCLES else None
sink = torch.zeros(4, dtype=torch.int32, device="cuda")
for name, mc in (("conv_only", at.ModelConfig(USE_TRANSFORMER=False)), ("default_tf", at.ModelConfig())):
    env = at.EnvConfig()
    big = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=prec, max_batch=B)
    hog = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=prec, max_batch=max(HOG, 64))
    w = synth_state_dict(big.get_weights(), 1)
    big.set_weights(w)
    hog.set_weights(w)
    base = pack_states(synth_states(256, env, seed=5), env)
    x_big = torch.from_numpy(np.tile(base, (-(-B // 256), 1))[:B]).cuda()
    x_hog = torch.from_numpy(base[:HOG].copy()).cuda()
    side = torch.cuda.Stream()

    def timed(steps, with_hog):
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if with_hog:
            with torch.cuda.stream(side):
                if HOG_CYCLES:  # HOG CTAs of the probe library's hog kernel, each a whole SM, for the whole timed region
                    _lib.check_probe(probe.at_probe_hog(HOG, int(HOG_CYCLES * steps), sink.data_ptr(), side.cuda_stream), "at_probe_hog")
                else:
                    for _ in range(steps * 40):
                        hog.engine.evaluate_packed(x_hog, hog.precision)
            time.sleep(0.002)  # let the hog get its SMs first
        ev0.record()
        for _ in range(steps):
            out = big.engine.evaluate_packed(x_big, big.precision)
        ev1.record()
        torch.cuda.synchronize()
        return ev0.elapsed_time(ev1) / steps, out

    timed(3, False)
    alone, ref = timed(20, False)
    shared, got = timed(20, True)
    same = all(torch.equal(a, b) for a, b in zip(ref[:2], got[:2]))
    print(f"{name} {prec} batch={B} hog={HOG} boards: alone {alone:.4f} ms/step, with the hog {shared:.4f} ms/step "
          f"(x{shared / alone:.3f}); results identical: {same}", flush=True)
