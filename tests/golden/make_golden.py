"""Generate tests/golden/*.npz by running the UNMODIFIED reference in this container.

    python tests/golden/make_golden.py

Imports `alphatriangle.features` and `alphatriangle.nn` from /root/reference (with
the import shims under oracle/refstubs standing in for the two compiled third-party
packages), feeds them the seeded states of alphatriangle_b200.synth and the seeded
weights of synth_state_dict, and records what the reference returns:

  packed   uint8 [B,80]   the states as PackedState records (so the GPU test needs no generator)
  grid     f32 [B,C,R,W]  extract_state_features(...)["grid"]
  other    f32 [B,D]      extract_state_features(...)["other_features"]
  policy_logits / value_logits   AlphaTriangleNet.forward on those features (fp32, CPU)
  probs    f32 [B,A]      NeuralNetwork.evaluate_batch policy dicts, densified
  values   f32 [B]        NeuralNetwork.evaluate_batch expected values

The reference's own test-suite holds no golden vectors for this path, so these
files are the pin for the oracle and for the CUDA kernels. /root/reference does not
exist on the GPU box; only the committed .npz files travel.
"""

from __future__ import annotations

import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, str(ROOT / "oracle" / "refstubs"))
os.environ.setdefault("NUMBA_CACHE_DIR", "/tmp/numba_cache")

import numpy as np  # noqa: E402
import torch  # noqa: E402

from golden_cases import CASES, case_configs, case_states, case_weights  # noqa: E402

from alphatriangle_b200.packing import pack_states  # noqa: E402


def main() -> None:
    from alphatriangle.config import ModelConfig as RefModelConfig
    from alphatriangle.config import TrainConfig as RefTrainConfig
    from alphatriangle.features import extract_state_features as ref_extract
    from alphatriangle.nn import NeuralNetwork as RefNN
    from trianglengin import EnvConfig as RefEnvConfig

    import alphatriangle_b200 as ours

    torch.set_num_threads(1)
    out_dir = Path(__file__).resolve().parent
    for name in CASES:
        mc_kwargs, env_kwargs, seed = case_configs(name)
        ref_mc = RefModelConfig(**mc_kwargs)
        ref_env = RefEnvConfig(**env_kwargs)
        our_env = ours.EnvConfig(**env_kwargs)
        states = case_states(name, our_env)

        nn = RefNN(ref_mc, ref_env, RefTrainConfig(COMPILE_MODEL=False, DEVICE="cpu", WORKER_DEVICE="cpu"), torch.device("cpu"))
        # our parameter container must carry exactly the reference's keys / shapes / dtypes
        mine = ours.AlphaTriangleNet(ours.ModelConfig(**mc_kwargs), our_env).state_dict()
        ref_sd = nn._orig_model.state_dict()
        assert list(mine.keys()) == list(ref_sd.keys()), (name, set(mine) ^ set(ref_sd))
        for k in ref_sd:
            assert mine[k].shape == ref_sd[k].shape and mine[k].dtype == ref_sd[k].dtype, (name, k)
        nn.set_weights(case_weights(name, ref_sd))

        feats = [ref_extract(s, ref_mc) for s in states]
        grid = np.stack([f["grid"] for f in feats])
        other = np.stack([f["other_features"] for f in feats])
        with torch.no_grad():
            pl, vl = nn.model(torch.from_numpy(grid), torch.from_numpy(other))
        results = nn.evaluate_batch(states)
        action_dim = nn.action_dim
        probs = np.zeros((len(states), action_dim), dtype=np.float32)
        values = np.zeros((len(states),), dtype=np.float32)
        for i, (pol, val) in enumerate(results):
            assert len(pol) == action_dim
            probs[i] = [pol[a] for a in range(action_dim)]
            values[i] = val
        single_pol, single_val = nn.evaluate_state(states[0])
        np.savez_compressed(
            out_dir / f"{name}.npz",
            packed=pack_states(states, our_env),
            grid=grid,
            other=other,
            policy_logits=pl.numpy(),
            value_logits=vl.numpy(),
            probs=probs,
            values=values,
            single_probs=np.array([single_pol[a] for a in range(action_dim)], dtype=np.float32),
            single_value=np.float32(single_val),
        )
        entropy = -(probs * np.log(np.maximum(probs, 1e-30))).sum(axis=1)
        print(f"{name}: B={len(states)} grid{grid.shape} other{other.shape} probs{probs.shape} "
              f"value[min,max]=[{values.min():.4f},{values.max():.4f}] keys={len(ref_sd)} policy entropy mean {entropy.mean():.3f} "
              f"nats, max prob > 0.3 on {(probs.max(axis=1) > 0.3).mean():.0%} of the states")


if __name__ == "__main__":
    main()
