"""at_engine_submit_host / at_engine_wait_host: two batches in flight give the bits of the blocking call."""
import numpy as np
import pytest
import torch

import alphatriangle_b200 as at
from alphatriangle_b200 import _lib
from alphatriangle_b200.packing import pack_states
from alphatriangle_b200.synth import synth_state_dict, synth_states

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_two_batches_in_flight_match_blocking_calls(precision):
    env, mc = at.EnvConfig(), at.ModelConfig(USE_TRANSFORMER=False)
    nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=precision, max_batch=512)
    nn.set_weights(synth_state_dict(nn.get_weights(), 11))
    nn._sync_engine()
    batches = [pack_states(synth_states(n, env, seed=100 + n), env) for n in (512, 300, 1, 512, 77)]
    want = []
    for b in batches:
        p, v, f = nn.engine.evaluate_host(b, nn.precision)
        want.append((p.copy(), v.copy(), f))
    got = [None] * len(batches)
    for i, b in enumerate(batches):
        if i >= 2:
            p, v, f = nn.engine.wait_host(i & 1)
            got[i - 2] = (p.copy(), v.copy(), f)
        nn.engine.submit_host(b, nn.precision, i & 1)
    for i in range(len(batches) - 2, len(batches)):
        p, v, f = nn.engine.wait_host(i & 1)
        got[i] = (p.copy(), v.copy(), f)
    for (p0, v0, f0), (p1, v1, f1) in zip(want, got):
        assert f0 == 0 and f1 == 0
        assert np.array_equal(p0, p1) and np.array_equal(v0, v1)


def test_slot_protocol_errors():
    env, mc = at.EnvConfig(), at.ModelConfig(USE_TRANSFORMER=False)
    nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision="bf16", max_batch=64)
    nn.set_weights(synth_state_dict(nn.get_weights(), 12))
    nn._sync_engine()
    b = pack_states(synth_states(8, env, seed=1), env)
    with pytest.raises(_lib.LibraryError, match="nothing was submitted"):
        nn.engine.wait_host(0)
    nn.engine.submit_host(b, nn.precision, 0)
    with pytest.raises(_lib.LibraryError, match="still in flight"):
        nn.engine.submit_host(b, nn.precision, 0)
    nn.engine.wait_host(0)
    with pytest.raises(ValueError, match="slot must be 0 or 1"):
        nn.engine.submit_host(b, nn.precision, 2)
    assert nn.engine.lib.at_engine_wait_host(nn.engine.handle, 5) != 0 and "slot must be 0 or 1" in _lib.last_error()
    with pytest.raises(ValueError):
        nn.engine.submit_host(pack_states(synth_states(65, env, seed=2), env), nn.precision, 1)


def test_evaluate_batch_with_more_states_than_max_batch():
    """network._evaluate_rows sends the chunks through the two host slots; results keep their order."""
    env, mc = at.EnvConfig(), at.ModelConfig(USE_TRANSFORMER=False)
    small = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision="fp32", max_batch=16)
    big = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision="fp32", max_batch=128)
    sd = synth_state_dict(small.get_weights(), 21)
    small.set_weights(sd)
    big.set_weights(sd)
    for n in (16, 17, 33, 50, 100):  # 1, 2, 3, 4 and 7 chunks
        states = synth_states(n, env, seed=300 + n)
        a, b = small.evaluate_batch(states), big.evaluate_batch(states)
        assert len(a) == len(b) == n
        for (da, va), (db, vb) in zip(a, b):
            assert abs(va - vb) <= 1e-5
            assert np.abs(np.array(list(da.values())) - np.array(list(db.values()))).max() <= 1e-6
