"""LeafBatcher in front of the real NeuralNetwork: coalesced calls return what direct calls return."""
import threading

import numpy as np
import pytest
import torch

import alphatriangle_b200 as at
from alphatriangle_b200.synth import synth_state_dict, synth_states

pytestmark = pytest.mark.gpu


def test_batched_results_equal_direct_results():
    env, mc = at.EnvConfig(), at.ModelConfig(USE_TRANSFORMER=False)
    nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision="bf16", max_batch=1024)
    nn.set_weights(synth_state_dict(nn.get_weights(), 3))
    pool = synth_states(16 * 12, env, seed=8)
    direct = nn.evaluate_batch(pool)
    got = [None] * 16
    with at.LeafBatcher(nn, max_batch=1024, max_wait_s=0.01) as b:
        def game(g, c):
            with c:
                got[g] = c.evaluate_batch(pool[g * 12 : (g + 1) * 12])
        threads = [threading.Thread(target=game, args=(g, b.client())) for g in range(16)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(60)
        assert b.calls < 16
    flat = [r for g in got for r in g]
    assert len(flat) == len(direct)
    for (d0, v0), (d1, v1) in zip(direct, flat):
        # a board's bits do not depend on where it sits in a launch; only the split-K factor differs with the batch size
        assert abs(v0 - v1) <= 5e-2
        a0, a1 = np.array(list(d0.values())), np.array(list(d1.values()))
        assert np.abs(a0 - a1).max() <= 1e-3 and abs(a1.sum() - 1.0) < 1e-5
