"""The drop-in boundary, written the way the reference's tests/nn/test_network.py is:
shapes, types, sum-to-one, weight round trip, NaN => NetworkEvaluationError."""

import numpy as np
import pytest
import torch
from golden_cases import case_configs
from gpu_util import load_case, network

import alphatriangle_b200 as at
from alphatriangle_b200 import _lib
from alphatriangle_b200.synth import synth_state_dict, synth_states

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nn_interface():
    return network("ref_test_cfg", "fp32")


@pytest.fixture
def game_state(nn_interface):
    return synth_states(1, nn_interface.env_config, seed=123)[0]


def test_network_initialization(nn_interface):
    env = nn_interface.env_config
    assert nn_interface.model is not None and nn_interface._orig_model is nn_interface.model
    assert nn_interface.device.type == "cuda"
    assert nn_interface.action_dim == env.NUM_SHAPE_SLOTS * env.ROWS * env.COLS
    assert nn_interface.num_atoms == 51 and nn_interface.support.shape == (51,)
    assert abs(nn_interface.delta_z - 0.4) < 1e-12


def test_state_to_tensors(nn_interface, game_state):
    mc, env = nn_interface.model_config, nn_interface.env_config
    grid, other = nn_interface._state_to_tensors(game_state)
    assert grid.shape == (1, mc.GRID_INPUT_CHANNELS, env.ROWS, env.COLS)
    assert other.shape == (1, mc.OTHER_NN_INPUT_FEATURES_DIM)
    assert grid.device == nn_interface.device and other.device == nn_interface.device


def test_batch_states_to_tensors(nn_interface, game_state):
    mc, env = nn_interface.model_config, nn_interface.env_config
    grid, other = nn_interface._batch_states_to_tensors([game_state.copy() for _ in range(3)])
    assert grid.shape == (3, mc.GRID_INPUT_CHANNELS, env.ROWS, env.COLS)
    assert other.shape == (3, mc.OTHER_NN_INPUT_FEATURES_DIM)
    g0, o0 = nn_interface._batch_states_to_tensors([])
    assert g0.shape == (0, mc.GRID_INPUT_CHANNELS, env.ROWS, env.COLS) and o0.shape == (0, mc.OTHER_NN_INPUT_FEATURES_DIM)


def test_evaluate_state(nn_interface, game_state):
    policy, value = nn_interface.evaluate_state(game_state)
    assert isinstance(policy, dict) and isinstance(value, float)
    assert len(policy) == nn_interface.action_dim
    assert all(isinstance(k, int) for k in policy) and all(isinstance(v, float) for v in policy.values())
    assert abs(sum(policy.values()) - 1.0) < 1e-5


def test_evaluate_batch(nn_interface, game_state):
    states = [game_state.copy() for _ in range(3)]
    results = nn_interface.evaluate_batch(states)
    assert isinstance(results, list) and len(results) == 3
    for policy, value in results:
        assert isinstance(policy, dict) and isinstance(value, float) and len(policy) == nn_interface.action_dim
        assert all(isinstance(k, int) for k in policy) and all(isinstance(v, float) for v in policy.values())
        assert abs(sum(policy.values()) - 1.0) < 1e-5
    assert nn_interface.evaluate_batch([]) == []


def test_evaluate_matches_reference_values(golden_dir):
    """evaluate_batch / evaluate_state dictionaries against what the reference's own methods returned."""
    from golden_cases import case_states

    mc, env, _seed, g = load_case("default_tf", golden_dir)
    nn = network("default_tf", "fp32")
    states = case_states("default_tf", env)
    res = nn.evaluate_batch(states)
    for i, (pol, val) in enumerate(res):
        assert np.abs(np.array([pol[a] for a in range(nn.action_dim)], dtype=np.float32) - g["probs"][i]).max() <= 1e-4
        assert abs(val - float(g["values"][i])) <= 1e-4
    pol, val = nn.evaluate_state(states[0])
    assert np.abs(np.array([pol[a] for a in range(nn.action_dim)], dtype=np.float32) - g["single_probs"]).max() <= 1e-4
    assert abs(val - float(g["single_value"])) <= 1e-4


def test_get_set_weights():
    mc, env, _seed = case_configs("ref_test_cfg")
    nn = at.NeuralNetwork(at.ModelConfig(**mc), at.EnvConfig(**env), at.TrainConfig(), torch.device("cuda:0"), precision="fp32", max_batch=64)
    initial = nn.get_weights()
    assert all(isinstance(v, torch.Tensor) and v.device == torch.device("cpu") for v in initial.values())
    state = synth_states(1, nn.env_config, seed=1)[0]
    before = nn.evaluate_state(state)
    modified = {k: (torch.zeros_like(v) if v.dtype.is_floating_point else v.clone()) for k, v in initial.items()}
    nn.set_weights(modified)
    after = nn.get_weights()
    for k in modified:
        if modified[k].dtype.is_floating_point:
            assert torch.allclose(after[k], modified[k], atol=1e-6)
        else:
            assert torch.equal(after[k], modified[k])
    # all-zero weights (running_var = 0 too): uniform policy, zero expected value -- the engine saw the new weights
    pol, val = nn.evaluate_state(state)
    assert max(abs(p - 1.0 / nn.action_dim) for p in pol.values()) < 1e-6 and abs(val) < 1e-5
    nn.set_weights(initial)
    again = nn.evaluate_state(state)
    assert again[1] == before[1] and again[0] == before[0]
    with pytest.raises(Exception):
        nn.set_weights({"nope": torch.zeros(1)})


def test_in_place_parameter_updates_are_picked_up():
    """The trainer steps `nn.model` in place; the next evaluation must use the new values."""
    mc, env, _seed = case_configs("ref_test_cfg")
    nn = at.NeuralNetwork(at.ModelConfig(**mc), at.EnvConfig(**env), None, torch.device("cuda:0"), precision="fp32", max_batch=8)
    nn.set_weights(synth_state_dict(nn.get_weights(), 11))
    state = synth_states(1, nn.env_config, seed=2)[0]
    v0 = nn.evaluate_state(state)[1]
    with torch.no_grad():
        last = [m for m in nn.model.value_head if isinstance(m, torch.nn.Linear)][-1]
        last.bias[0] += 25.0
    v1 = nn.evaluate_state(state)[1]
    assert v1 < v0 - 1.0  # atom 0 is VALUE_MIN


def test_evaluate_with_nan_features(nn_interface, game_state):
    grid, other = nn_interface._batch_states_to_tensors([game_state.copy(), game_state.copy()])
    other = other.clone()
    other[0, 0] = float("nan")
    with pytest.raises(at.NetworkEvaluationError, match="Non-finite values found"):
        nn_interface.evaluate_features(grid, other)
    grid = grid.clone()
    grid[1, 0, 0, 0] = float("inf")
    with pytest.raises(at.NetworkEvaluationError, match="Non-finite values found in batched input grid_tensor"):
        nn_interface.evaluate_features(grid, other)


def test_nonfinite_weights_raise(game_state):
    mc, env, _seed = case_configs("ref_test_cfg")
    nn = at.NeuralNetwork(at.ModelConfig(**mc), at.EnvConfig(**env), None, torch.device("cuda:0"), precision="fp32", max_batch=8)
    w = nn.get_weights()
    w["policy_head.3.bias"] = torch.full_like(w["policy_head.3.bias"], float("nan"))
    nn.set_weights(w)
    with pytest.raises(at.NetworkEvaluationError, match="Non-finite policy_logits"):
        nn.evaluate_state(game_state)
    with pytest.raises(at.NetworkEvaluationError, match="Batch evaluation failed"):
        nn.evaluate_batch([game_state])


def test_zero_policy_mass_falls_back_to_valid_actions(nn_interface, game_state):
    row = np.zeros(nn_interface.action_dim, dtype=np.float32)
    game_state._valid = [1, 4]
    out = nn_interface._finish_policy(row, game_state, "t")
    assert out[1] == 0.5 and out[4] == 0.5 and out.sum() == 1.0
    game_state._valid = []
    with pytest.raises(at.NetworkEvaluationError, match="no valid actions"):
        nn_interface._finish_policy(row, game_state, "t")


def test_engine_rejects_bad_calls(nn_interface):
    eng = nn_interface.engine
    with pytest.raises(ValueError):
        eng.evaluate_packed(torch.zeros((2, 79), dtype=torch.uint8, device="cuda"), _lib.PREC_FP32)
    with pytest.raises(_lib.LibraryError, match="floats"):
        eng.load_image(torch.zeros(3, device="cuda"))
    p, v, f = eng.evaluate_packed(torch.zeros((0, 80), dtype=torch.uint8, device="cuda"), _lib.PREC_FP32)
    assert p.shape == (0, nn_interface.action_dim) and v.shape == (0,)
