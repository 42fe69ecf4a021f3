"""Weight production and publication on the GPU (first slice).

The reference's trainer steps `nn.model` with a torch optimiser (alphatriangle/rl/core/trainer.py:204-310) and the
training loop pushes the result to the self-play workers while they keep playing (training/loop.py:271-287,
training/worker_manager.py:169-209). Here the parameters live inside the engine's weight image, a publication is one
collective over that buffer on a side stream, and evaluations switch sets only between calls."""

import os
import socket

import numpy as np
import pytest
import torch

import alphatriangle_b200 as at
from alphatriangle_b200 import _lib
from alphatriangle_b200.packing import pack_states
from alphatriangle_b200.synth import synth_state_dict, synth_states

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _net(mc, precision, seed, max_batch=512):
    nn = at.NeuralNetwork(mc, at.EnvConfig(), None, DEV, precision=precision, max_batch=max_batch)
    nn.set_weights(synth_state_dict(nn.get_weights(), seed, sharpness=1.0))
    return nn


def _train_step(nn, opt, batch=32, seed=0):
    """Trainer-side forward / backward in train mode (batch-statistics BatchNorm, dropout) and one optimiser step on
    `nn.model`, exactly what the reference's Trainer.train_step does with it."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    env, mc = nn.env_config, nn.model_config
    grid = torch.randn((batch, mc.GRID_INPUT_CHANNELS, env.ROWS, env.COLS), generator=g).to(DEV)
    other = torch.randn((batch, mc.OTHER_NN_INPUT_FEATURES_DIM), generator=g).to(DEV)
    target_p = torch.softmax(torch.randn((batch, nn.action_dim), generator=g), dim=1).to(DEV)
    target_v = torch.softmax(torch.randn((batch, nn.num_atoms), generator=g), dim=1).to(DEV)
    nn.model.train()
    pl, vl = nn.model(grid, other)
    loss = -(target_p * torch.log_softmax(pl, 1)).sum(1).mean() - (target_v * torch.log_softmax(vl, 1)).sum(1).mean()
    opt.zero_grad(set_to_none=True)
    loss.backward()
    opt.step()
    nn.model.eval()
    return float(loss)


@pytest.mark.parametrize("tf", [False, True])
def test_optimiser_step_writes_the_image_and_the_engine_follows(tf):
    mc = at.ModelConfig(USE_TRANSFORMER=tf)
    nn = _net(mc, "fp32", 5, max_batch=64)
    assert nn.flat.aliased()
    packed = torch.from_numpy(pack_states(synth_states(48, nn.env_config, seed=8), nn.env_config)).to(DEV)
    p0, v0 = [t.clone() for t in nn.evaluate_packed(packed)]
    tracked0 = int(nn.get_weights()["conv_body.0.1.num_batches_tracked"])
    opt = torch.optim.AdamW(nn.model.parameters(), lr=3e-3)
    image_before = nn.flat.image.clone()
    losses = [_train_step(nn, opt, seed=s) for s in range(3)]
    assert all(np.isfinite(losses)) and nn.flat.aliased()
    assert not torch.equal(image_before, nn.flat.image)  # the optimiser wrote the engine's input buffer
    p1, v1 = [t.clone() for t in nn.evaluate_packed(packed)]  # version counters moved -> re-folded before this call
    assert not torch.equal(p0, p1)
    # a fresh network given the same state_dict (BatchNorm running statistics included) returns the same bits
    twin = at.NeuralNetwork(mc, at.EnvConfig(), None, DEV, precision="fp32", max_batch=64)
    twin.set_weights(nn.get_weights())
    p2, v2 = twin.evaluate_packed(packed)
    assert torch.equal(p1, p2) and torch.equal(v1, v2)
    assert int(nn.get_weights()["conv_body.0.1.num_batches_tracked"]) == tracked0 + 3


@pytest.mark.parametrize("precision,tf", [("fp32", False), ("bf16", False), ("bf16", True)])
def test_evaluations_racing_a_publication_see_one_weight_set(precision, tf):
    """Evaluations run back to back on the main stream while new weights are folded into the inactive set on a side
    stream: every result equals the old weights' bits or the new weights' bits, never a mixture, and the switch
    happens."""
    mc = at.ModelConfig(USE_TRANSFORMER=tf)
    nn = _net(mc, precision, 21)
    eng = nn.engine
    n = 512 if precision != "fp32" else 96
    packed = torch.from_numpy(pack_states(synth_states(n, nn.env_config, seed=77), nn.env_config)).to(DEV)
    image = {0: nn.flat.image.clone()}
    other = synth_state_dict(nn.get_weights(), 22, sharpness=1.0)
    image[1] = eng.build_image(other)
    want = {}
    for k in (0, 1):
        eng.load_image(image[k], 100 + k)
        p, v, _f = eng.evaluate_packed(packed, nn.precision)
        want[k] = (p.clone(), v.clone())
    assert not torch.equal(want[0][0], want[1][0])
    torch.cuda.synchronize()
    side = torch.cuda.Stream(device=DEV)
    current, seen_switch, results = 1, 0, []
    for it in range(60):
        if it % 10 == 3:  # publish the other set on the side stream, evaluations keep going
            current ^= 1
            with torch.cuda.stream(side):
                eng.publish_image(image[current], 200 + it)
        p, v, _f = eng.evaluate_packed(packed, nn.precision)
        results.append((p.clone(), v.clone(), eng.weight_version()))
        # a search consumes each batch's results before it can build the next one: the host waits for ITS stream (never
        # for the side stream). The switch-over decision is taken on the host when a call is queued, so a host that ran
        # far ahead of the GPU would simply see it later.
        torch.cuda.current_stream(DEV).synchronize()
    torch.cuda.synchronize()
    versions = [r[2] for r in results]
    for p, v, ver in results:
        k = 0 if torch.equal(p, want[0][0]) else 1
        assert torch.equal(p, want[k][0]) and torch.equal(v, want[k][1]), "an evaluation mixed two weight sets"
    seen_switch = sum(1 for a, b in zip(versions, versions[1:]) if a != b)
    assert seen_switch >= 5, versions  # six publications, each taken over within the ten calls that follow


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_publisher_one_collective_on_a_side_stream():
    """Publisher.publish through NCCL (a one-rank group here; the N-rank case runs in bench.py --gpus N): the header tail
    carries the version and the integer buffers, no host synchronisation in between."""
    import torch.distributed as dist

    from alphatriangle_b200.dist import Publisher, received_version

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(_free_port())
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=DEV)
    try:
        nn = _net(at.ModelConfig(USE_TRANSFORMER=False), "bf16", 31)
        packed = torch.from_numpy(pack_states(synth_states(256, nn.env_config, seed=5), nn.env_config)).to(DEV)
        p0, _v0 = [t.clone() for t in nn.evaluate_packed(packed)]
        pub = Publisher(nn)
        with torch.no_grad():
            last = [m for m in nn.model.policy_head if isinstance(m, torch.nn.Linear)][-1]
            last.weight.mul_(0.5)
        pub.publish(src=0, version=4242)
        pub.wait()
        assert received_version(nn.flat) == (4242, nn.flat.image_floats)
        p1, _v1 = nn.evaluate_packed(packed)
        assert nn.engine.weight_version() == 4242 and not torch.equal(p0, p1)
    finally:
        dist.destroy_process_group()
