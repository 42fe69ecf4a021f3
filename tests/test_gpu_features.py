"""Feature kernel parity through the C ABI: bit-exact against the reference's recorded
outputs (golden) and against the oracle on fresh seeded states, plus size-independent
properties at BASELINE batch sizes."""

import numpy as np
import pytest
import torch
from golden_cases import CASES
from gpu_util import bits, load_case

import alphatriangle_b200 as at
from alphatriangle_b200.features import extract_features_batch, extract_state_features
from alphatriangle_b200.packing import pack_states
from alphatriangle_b200.synth import edge_states, synth_states
from oracle import features_ref

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", CASES)
def test_features_bit_exact_vs_reference_golden(name, golden_dir):
    mc, env, _seed, g = load_case(name, golden_dir)
    from alphatriangle_b200.features import _run

    grid, other = _run(g["packed"], env, mc.GRID_INPUT_CHANNELS, mc.OTHER_NN_INPUT_FEATURES_DIM, torch.device("cuda:0"))
    assert np.array_equal(bits(grid), bits(g["grid"]))
    assert np.array_equal(bits(other), bits(g["other"]))


def test_features_bit_exact_vs_oracle_2048_states():
    env, mc = at.EnvConfig(), at.ModelConfig()
    states = synth_states(2048, env, seed=99) + edge_states(env)
    feats = extract_features_batch(states, mc)
    grid_o, other_o = features_ref.extract_batch(states, mc)
    assert np.array_equal(bits(np.stack([f["grid"] for f in feats])), bits(grid_o))
    assert np.array_equal(bits(np.stack([f["other_features"] for f in feats])), bits(other_o))


def test_single_state_api_and_empty_batch():
    env, mc = at.EnvConfig(), at.ModelConfig()
    s = synth_states(1, env, seed=5)[0]
    f = extract_state_features(s, mc)
    o = features_ref.extract_state_features(s, mc)
    assert f["grid"].shape == (1, 8, 15) and f["grid"].dtype == np.float32
    assert f["other_features"].shape == (30,) and f["other_features"].dtype == np.float32
    assert np.array_equal(bits(f["grid"]), bits(o["grid"])) and np.array_equal(bits(f["other_features"]), bits(o["other_features"]))
    assert extract_features_batch([], mc) == []


@pytest.mark.parametrize("batch", [4096, 65536])
def test_features_full_size_properties(batch):
    """At BASELINE sizes: tiling a small seeded batch must reproduce the small batch's rows exactly
    (every block / tail path gives the same answer), and the value ranges the domain guarantees hold."""
    env, mc = at.EnvConfig(), at.ModelConfig()
    from alphatriangle_b200.features import _run

    base_states = synth_states(509, env, seed=3) + edge_states(env)  # 521 rows: not a multiple of the block's 16 states
    base = pack_states(base_states, env)
    reps = -(-batch // base.shape[0])
    packed = np.tile(base, (reps, 1))[:batch]
    grid, other = _run(packed, env, 1, 30, torch.device("cuda:0"))
    g0, o0 = features_ref.extract_batch(base_states, mc)
    idx = np.arange(batch) % base.shape[0]
    assert np.array_equal(bits(grid), bits(g0)[idx])
    assert np.array_equal(bits(other), bits(o0)[idx])
    assert set(np.unique(grid)) <= {-1.0, 0.0, 1.0}
    assert other[:, :24].min() >= 0.0 and other[:, :24].max() <= 1.0
    assert np.abs(other[:, 24]).max() <= 5.0 and np.isfinite(other).all()
