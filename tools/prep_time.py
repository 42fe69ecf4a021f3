"""Device time of one weight load (at_engine_load_weights: the two weight-preparation launches) per model."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import alphatriangle_b200 as at  # noqa: E402
from alphatriangle_b200.synth import synth_state_dict  # noqa: E402

for tf in (False, True):
    nn = at.NeuralNetwork(at.ModelConfig(USE_TRANSFORMER=tf), at.EnvConfig(), None, torch.device("cuda:0"), max_batch=256)
    nn.set_weights(synth_state_dict(nn.get_weights(), 1))
    ts = []
    for i in range(8):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        nn.engine.load_image(nn.flat.image, 10 + i)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    print(f"transformer={tf}: load_image device ms min {min(ts):.4f} median {sorted(ts)[4]:.4f}")
