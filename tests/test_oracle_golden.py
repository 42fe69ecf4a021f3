"""The oracle against the reference's recorded outputs (tests/golden/*.npz, made by
make_golden.py from the unmodified reference). Features must be bit-exact; the
restated forward must agree with the reference's module forward to fp32 round-off."""

import numpy as np
import pytest
import torch
from golden_cases import CASES, case_configs, case_states, case_weights

import alphatriangle_b200 as at
from alphatriangle_b200.packing import pack_states
from oracle import features_ref, model_ref


def _load(name, golden_dir):
    mc_kwargs, env_kwargs, seed = case_configs(name)
    mc = at.ModelConfig(**mc_kwargs)
    env = at.EnvConfig(**env_kwargs)
    return mc, env, seed, np.load(golden_dir / f"{name}.npz")


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


@pytest.mark.parametrize("name", CASES)
def test_generator_and_packer_are_stable(name, golden_dir):
    mc, env, seed, g = _load(name, golden_dir)
    packed = pack_states(case_states(name, env), env)
    assert packed.dtype == np.uint8 and packed.shape == g["packed"].shape
    assert np.array_equal(packed, g["packed"])


@pytest.mark.parametrize("name", CASES)
def test_feature_oracle_bit_exact(name, golden_dir):
    mc, env, seed, g = _load(name, golden_dir)
    grid, other = features_ref.extract_batch(case_states(name, env), mc)
    assert grid.dtype == np.float32 and other.dtype == np.float32
    assert np.array_equal(_bits(grid), _bits(g["grid"]))
    assert np.array_equal(_bits(other), _bits(g["other"]))  # includes the sign of zero


@pytest.mark.parametrize("name", CASES)
def test_forward_oracle_matches_reference(name, golden_dir):
    mc, env, seed, g = _load(name, golden_dir)
    torch.set_num_threads(1)
    sd = case_weights(name, at.AlphaTriangleNet(mc, env).state_dict())
    pl, vl = model_ref.forward(sd, mc, torch.from_numpy(g["grid"]), torch.from_numpy(g["other"]))
    # fp32 summation order differs between the module tree and the functional restatement: the round-off scales with
    # the logits, which are O(0.1) with the small weights and O(10-50) with the trained-like ones
    scale = max(1.0, float(np.abs(g["policy_logits"]).max()), float(np.abs(g["value_logits"]).max()))
    assert np.abs(pl.numpy() - g["policy_logits"]).max() < 2e-5 * scale
    assert np.abs(vl.numpy() - g["value_logits"]).max() < 2e-5 * scale
    probs, values = model_ref.evaluate_tensors(sd, mc, g["grid"], g["other"])
    assert np.abs(probs - g["probs"]).max() < 1e-6 * scale
    assert np.abs(values - g["values"]).max() < 1e-5 * scale


@pytest.mark.parametrize("name", ["ref_test_cfg", "odd_gelu"])
def test_evaluate_batch_oracle_structure(name, golden_dir):
    mc, env, seed, g = _load(name, golden_dir)
    sd = case_weights(name, at.AlphaTriangleNet(mc, env).state_dict())
    states = case_states(name, env)
    res = model_ref.evaluate_batch(sd, mc, states)
    assert len(res) == len(states)
    a = env.NUM_SHAPE_SLOTS * env.ROWS * env.COLS
    for i, (pol, val) in enumerate(res):
        assert len(pol) == a and all(isinstance(k, int) for k in pol) and isinstance(val, float)
        assert abs(sum(pol.values()) - 1.0) < 1e-5
        assert abs(val - float(g["values"][i])) < 1e-5
    assert model_ref.evaluate_batch(sd, mc, []) == []
