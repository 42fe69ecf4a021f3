"""at_engine_evaluate_packed_static: evaluation of small batches with no device allocation.

Some hosts cannot allocate device memory once they are running: a self-play process sharing its GPU with the
trainer (a cudaMalloc in the search loop synchronises the device and can fail when the trainer owns the rest of
the memory), or an evaluator that keeps one engine per model version and cannot afford private scratch for each.
For them the library keeps the buffers its kernels hand to one another in a single static device array (6 MiB,
shared by every engine on the GPU), and this entry point evaluates up to 128 states through it.

The tests call it through the C ABI on an engine created with max_batch = 1 -- so the engine's own buffers could
not possibly hold the batch -- twice in a row with different batches (the second smaller, so it runs on top of the
first call's left-overs), for both BASELINE models, and compare with at_engine_evaluate_packed on a normally sized
engine. Same kernels, same split-K factors: the tolerance is the one tests/test_gpu_bf16.py uses for re-batched runs.
A second test checks the claim itself: cudaMemGetInfo and torch's allocator do not move across a series of calls in
which two models share the array. File name sorts last on purpose so a `-x` run has already been through every other
test.
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


def _pair_of_networks(case):
    """(normally sized engine, engine that owns scratch for ONE state only), same seeded weights."""
    mc_kwargs, env_kwargs, seed = case_configs(case)
    mc, env = at.ModelConfig(**mc_kwargs), at.EnvConfig(**env_kwargs)
    dev = torch.device("cuda:0")
    regular = at.NeuralNetwork(mc, env, None, dev, precision="bf16", max_batch=256)
    regular.set_weights(synth_state_dict(regular.get_weights(), seed))
    tiny = at.NeuralNetwork(mc, env, None, dev, precision="bf16", max_batch=1)
    tiny.set_weights(regular.get_weights())
    tiny._sync_engine()
    return env, regular, tiny


@pytest.mark.parametrize("case", ["conv_only", "default_tf"])  # tower_kernel / tower_tf_kernel through the static array
def test_static_scratch_entry_point_matches_the_allocating_one(case):
    env, regular, tiny = _pair_of_networks(case)
    dev = regular.device
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

    # beyond the static array's capacity: refused with a message, nothing queued, outputs untouched
    too_many = torch.from_numpy(pack_states((states * 2)[:129], env)).to(dev)
    rc, probs, value, _ = _static_eval(tiny, too_many)
    assert rc == _lib.AT_E_INVALID and "128" in _lib.last_error()
    torch.cuda.synchronize()
    assert float(probs.min().item()) == -1.0 and float(value.min().item()) == -1.0
    # empty batch is a no-op
    rc, _, _, _ = _static_eval(tiny, torch.zeros((0, 80), dtype=torch.uint8, device=dev))
    assert rc == 0


def test_static_entry_point_allocates_nothing_and_is_shared_between_models():
    """What the entry point is for: once the first call on the device has happened, further calls leave the device's free
    memory (cudaMemGetInfo) and torch's allocator exactly where they were -- also when engines of two different models
    take turns on the one array, each call running over the other model's left-overs."""
    env, reg_a, tiny_a = _pair_of_networks("conv_only")
    _env, reg_b, tiny_b = _pair_of_networks("default_tf")
    dev = reg_a.device
    states = synth_states(128, env, seed=99)
    packed = torch.from_numpy(pack_states(states, env)).to(dev)
    want = {}
    for name, reg in (("a", reg_a), ("b", reg_b)):
        p, v, _f = reg.engine.evaluate_packed(packed, _lib.PREC_BF16)
        want[name] = (p.clone(), v.clone())
    eng = {"a": tiny_a.engine, "b": tiny_b.engine}
    outs = {k: (torch.empty((128, e.action_dim), dtype=torch.float32, device=dev), torch.empty((128,), dtype=torch.float32, device=dev),
                torch.zeros(4, dtype=torch.int32, device=dev)) for k, e in eng.items()}
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    def call(k, n):
        p, v, f = outs[k]
        return eng[k].lib.at_engine_evaluate_packed_static(eng[k].handle, C.c_void_p(packed.data_ptr()), n, C.c_void_p(p.data_ptr()),
                                                           C.c_void_p(v.data_ptr()), C.c_void_p(f.data_ptr()), stream)

    assert call("a", 128) == 0 and call("b", 128) == 0, _lib.last_error()  # first calls: address look-up, lazy module load
    torch.cuda.synchronize()
    free0, _total = torch.cuda.mem_get_info(dev)
    torch_alloc0, torch_reserved0 = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)
    for n in (128, 5, 64, 128):
        for k in ("a", "b"):
            assert call(k, n) == 0, _lib.last_error()
    torch.cuda.synchronize()
    free1, _total = torch.cuda.mem_get_info(dev)
    assert free1 == free0, f"device free memory moved by {free0 - free1} bytes across eight static calls"
    assert torch.cuda.memory_allocated(dev) == torch_alloc0 and torch.cuda.memory_reserved(dev) == torch_reserved0
    for k in ("a", "b"):
        p, v, f = outs[k]
        assert int(f[0].item()) == 0
        assert np.abs(p.cpu().numpy() - want[k][0].cpu().numpy()).max() <= PROBS_TOL
        assert np.abs(v.cpu().numpy() - want[k][1].cpu().numpy()).max() <= VALUE_TOL
