"""at_engine_evaluate_packed_static: evaluation of small batches with no device allocation.

Some hosts cannot allocate device memory once they are running: a self-play process sharing its GPU with the
trainer (a cudaMalloc in the search loop synchronises the device and can fail when the trainer owns the rest of
the memory), or an evaluator that keeps one engine per model version and cannot afford private scratch for each.
For them the library keeps the buffers its kernels hand to one another in a single static device array (6 MiB,
shared by every engine on the GPU), and this entry point evaluates up to 128 states through it.

The test calls it through the C ABI on an engine created with max_batch = 1 -- so the engine's own buffers could
not possibly hold the batch -- twice in a row with different batches (the second smaller, so it runs on top of the
first call's left-overs), and compares with at_engine_evaluate_packed on a normally sized engine. Same kernels, same
split-K factors: the tolerance is the one tests/test_gpu_bf16.py uses for re-batched runs. File name sorts last
on purpose so a `-x` run has already been through every other test.
"""

import ctypes as C

import numpy as np
import pytest
import torch
from golden_cases import case_configs

import alphatriangle_b200 as at
from alphatriangle_b200 import _lib
from alphatriangle_b200.packing import pack_states
from alphatriangle_b200.synth import edge_states, synth_state_dict, synth_states

pytestmark = pytest.mark.gpu

PROBS_TOL = 1e-6
VALUE_TOL = 1e-4


def _static_eval(nn, packed: torch.Tensor):
    eng = nn.engine
    b = packed.shape[0]
    probs = torch.full((b, eng.action_dim), -1.0, dtype=torch.float32, device=eng.device)
    value = torch.full((b,), -1.0, dtype=torch.float32, device=eng.device)
    flags = torch.zeros(4, dtype=torch.int32, device=eng.device)
    rc = eng.lib.at_engine_evaluate_packed_static(
        eng.handle, C.c_void_p(packed.data_ptr()), b, C.c_void_p(probs.data_ptr()), C.c_void_p(value.data_ptr()),
        C.c_void_p(flags.data_ptr()), C.c_void_p(torch.cuda.current_stream(eng.device).cuda_stream),
    )
    return rc, probs, value, flags


def test_static_scratch_entry_point_matches_the_allocating_one():
    mc_kwargs, env_kwargs, seed = case_configs("conv_only")
    mc, env = at.ModelConfig(**mc_kwargs), at.EnvConfig(**env_kwargs)
    dev = torch.device("cuda:0")
    regular = at.NeuralNetwork(mc, env, None, dev, precision="bf16", max_batch=256)
    regular.set_weights(synth_state_dict(regular.get_weights(), seed))
    tiny = at.NeuralNetwork(mc, env, None, dev, precision="bf16", max_batch=1)  # owns scratch for ONE state only
    tiny.set_weights(regular.get_weights())
    tiny._sync_engine()

    states = synth_states(116, env, seed=31337) + edge_states(env)  # 128 states: the limit
    for n in (128, 37):  # second call: smaller batch on top of the first call's contents
        packed = torch.from_numpy(pack_states(states[:n], env)).to(dev)
        ref_p, ref_v, ref_f = regular.engine.evaluate_packed(packed, _lib.PREC_BF16)
        rc, probs, value, flags = _static_eval(tiny, packed)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        assert int(flags[0].item()) == 0 and int(ref_f[0].item()) == 0
        assert np.abs(probs.cpu().numpy() - ref_p.cpu().numpy()).max() <= PROBS_TOL
        assert np.abs(value.cpu().numpy() - ref_v.cpu().numpy()).max() <= VALUE_TOL
        assert np.abs(probs.cpu().numpy().sum(axis=1) - 1.0).max() < 1e-5

    # beyond the static array's capacity: refused with a message, outputs untouched
    too_many = torch.from_numpy(pack_states((states * 2)[:129], env)).to(dev)
    rc, probs, value, _ = _static_eval(tiny, too_many)
    assert rc == _lib.AT_E_INVALID and "128" in _lib.last_error()
    torch.cuda.synchronize()
    assert float(probs.min().item()) == -1.0 and float(value.min().item()) == -1.0
    # empty batch is a no-op
    rc, _, _, _ = _static_eval(tiny, torch.zeros((0, 80), dtype=torch.uint8, device=dev))
    assert rc == 0
