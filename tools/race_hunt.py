"""Hunt for run-to-run differences: evaluate a batch of repeated boards many times and report every board whose result
differs from the first copy of the same board (inside one launch) or from the first launch.

    python tools/race_hunt.py [launches] [batch] [flush]     # flush=1 writes a 512 MiB buffer between launches (cold L2)
"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "tests"), str(ROOT / "tests" / "golden")]
from gpu_util import load_case, network  # noqa: E402

launches = int(sys.argv[1]) if len(sys.argv) > 1 else 200
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
flush = int(sys.argv[3]) if len(sys.argv) > 3 else 0
golden = ROOT / "tests" / "golden"
junk = torch.empty(512 << 20, dtype=torch.uint8, device="cuda") if flush else None

import alphatriangle_b200 as at  # noqa: E402
from alphatriangle_b200.packing import pack_states  # noqa: E402
from alphatriangle_b200.synth import synth_state_dict, synth_states  # noqa: E402

SHAPES = {  # model shapes beyond the golden cases (same ones tests/test_gpu_bf16.py checks against the oracle)
    "tf_1layer_ff128_gelu": dict(TRANSFORMER_LAYERS=1, TRANSFORMER_FC_DIM=128, ACTIVATION_FUNCTION="GELU"),
    "tf_3layers_ff512": dict(TRANSFORMER_LAYERS=3, TRANSFORMER_FC_DIM=512),
    "tf_cuda_core_heads": dict(FC_DIMS_SHARED=[96], POLICY_HEAD_DIMS=[48], VALUE_HEAD_DIMS=[24]),
    "conv_unfused_tail": dict(USE_TRANSFORMER=False, FC_DIMS_SHARED=[128, 64], POLICY_HEAD_DIMS=[64, 32], VALUE_HEAD_DIMS=[]),
}
cases = sys.argv[4].split(",") if len(sys.argv) > 4 else ["default_tf", "conv_only"]

for case in cases:
    for fmt in ("f16", "bf16"):
        if case in SHAPES:
            mc, env = at.ModelConfig(**SHAPES[case]), at.EnvConfig()
            nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=fmt, max_batch=max(batch, 4096))
            nn.set_weights(synth_state_dict(nn.get_weights(), 91, sharpness=1.0))
            base = pack_states(synth_states(48, env, seed=777), env)
        else:
            _mc, _env, _seed, g = load_case(case, golden)
            nn = network(case, fmt, max_batch=max(batch, 4096))
            base = g["packed"]
        n = base.shape[0]
        packed = torch.from_numpy(np.tile(base, (-(-batch // n), 1))[:batch]).cuda()
        idx = torch.arange(batch, device="cuda") % n
        first = None
        bad_launches = 0
        for it in range(launches):
            if junk is not None:
                junk.fill_(it & 0xFF)
            probs, value = nn.evaluate_packed(packed)
            if first is None:
                first = (probs.clone(), value.clone())
            inside = (probs != probs[:n][idx]).any(dim=1) | (value != value[:n][idx])
            across = (probs != first[0]).any(dim=1) | (value != first[1])
            if inside.any() or across.any():
                bad_launches += 1
                rows = torch.nonzero(inside | across).flatten().cpu().numpy()
                dp = (probs - probs[:n][idx]).abs().max(dim=1).values.cpu().numpy()
                dv = (value - value[:n][idx]).abs().cpu().numpy()
                if bad_launches <= 6:
                    print(f"{case}/{fmt} launch {it}: {len(rows)} rows differ: {rows[:24].tolist()} "
                          f"max|dp|={dp.max():.3e} max|dv|={dv.max():.3e} inside={int(inside.sum())} across={int(across.sum())}",
                          flush=True)
        print(f"{case}/{fmt} batch={batch} flush={flush}: {bad_launches} of {launches} launches differ", flush=True)
