"""fp32 CUDA path parity (north_star: max-abs 1e-4 on policy and value against the
reference's fp32 path). The comparison target is what the unmodified reference returned
(tests/golden), with the oracle as a second witness on fresh states."""

import numpy as np
import pytest
import torch
from golden_cases import CASES
from gpu_util import load_case, network

import alphatriangle_b200 as at
from alphatriangle_b200.packing import pack_states
from alphatriangle_b200.synth import synth_states
from oracle import features_ref, model_ref

pytestmark = pytest.mark.gpu

TOL = 1e-4  # north_star tolerance for the fp32 path, policy probabilities and expected value


@pytest.mark.parametrize("name", CASES)
def test_evaluate_packed_matches_reference_golden(name, golden_dir):
    mc, env, _seed, g = load_case(name, golden_dir)
    nn = network(name, "fp32")
    probs, value = nn.evaluate_packed(torch.from_numpy(g["packed"]).cuda())
    probs, value = probs.cpu().numpy(), value.cpu().numpy()
    assert probs.shape == g["probs"].shape and value.shape == g["values"].shape
    assert np.abs(probs - g["probs"]).max() <= TOL
    assert np.abs(value - g["values"]).max() <= TOL
    assert np.abs(probs.sum(axis=1) - 1.0).max() < 1e-5


@pytest.mark.parametrize("name", ["default_tf", "conv_only", "odd_gelu"])
def test_forward_on_prepared_features_matches_reference(name, golden_dir):
    mc, env, _seed, g = load_case(name, golden_dir)
    nn = network(name, "fp32")
    probs, value = nn.evaluate_features(torch.from_numpy(g["grid"]), torch.from_numpy(g["other"]))
    assert np.abs(probs.cpu().numpy() - g["probs"]).max() <= TOL
    assert np.abs(value.cpu().numpy() - g["values"]).max() <= TOL


@pytest.mark.parametrize("name,batch", [("conv_only", 300), ("default_tf", 257)])
def test_fresh_states_match_oracle(name, batch, golden_dir):
    """Batches that are not tile multiples, against the oracle run on the same states and weights."""
    mc, env, _seed, _g = load_case(name, golden_dir)
    nn = network(name, "fp32")
    states = synth_states(batch, env, seed=4242)
    sd = {k: v for k, v in nn.get_weights().items()}
    grid, other = features_ref.extract_batch(states, mc)
    torch.set_num_threads(8)
    p_ref, v_ref = model_ref.evaluate_tensors(sd, mc, grid, other)
    probs, value = nn.evaluate_packed(torch.from_numpy(pack_states(states, env)).cuda())
    assert np.abs(probs.cpu().numpy() - p_ref).max() <= TOL
    assert np.abs(value.cpu().numpy() - v_ref).max() <= TOL


def test_chunked_large_batch_is_consistent(golden_dir):
    """A batch larger than the engine's internal slice (2048 boards) equals the same rows evaluated alone."""
    mc, env, _seed, g = load_case("conv_only", golden_dir)
    nn = network("conv_only", "fp32", max_batch=4096)
    base = g["packed"]
    reps = -(-4096 // base.shape[0])
    packed = np.tile(base, (reps, 1))[:4096]
    probs, value = nn.evaluate_packed(torch.from_numpy(packed).cuda())
    p0, v0 = nn.evaluate_packed(torch.from_numpy(base).cuda())
    idx = np.arange(4096) % base.shape[0]
    assert np.abs(probs.cpu().numpy() - p0.cpu().numpy()[idx]).max() <= 1e-6
    assert np.abs(value.cpu().numpy() - v0.cpu().numpy()[idx]).max() <= 1e-5
