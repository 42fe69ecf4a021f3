"""world_size-2 gloo run of the weight-publication host logic (CPU tensors, no GPU)."""

import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import alphatriangle_b200 as at
from alphatriangle_b200 import _lib
from alphatriangle_b200.dist import broadcast_flat, received_version
from alphatriangle_b200.flat import FlatWeights
from alphatriangle_b200.synth import synth_state_dict

SMALL = dict(CONV_FILTERS=[8, 16], CONV_KERNEL_SIZES=[3, 3], CONV_STRIDES=[1, 1], CONV_PADDING=[1, 1], NUM_RESIDUAL_BLOCKS=1,
             RESIDUAL_BLOCK_FILTERS=16, TRANSFORMER_DIM=16, TRANSFORMER_HEADS=2, TRANSFORMER_LAYERS=1, TRANSFORMER_FC_DIM=32)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _flat(model, mc, env):
    table = _lib.weight_table(mc, env)
    floats = max(o + n for _k, o, n in table)
    return FlatWeights(model, table, floats, torch.device("cpu"))


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mc, env = at.ModelConfig(**SMALL), at.EnvConfig()
        torch.manual_seed(100 + rank)  # ranks start with different weights
        model = at.AlphaTriangleNet(mc, env)
        flat = _flat(model, mc, env)
        if rank == 0:
            model.load_state_dict(synth_state_dict(model.state_dict(), 42))  # in place: lands in the flat image
            for v in model.state_dict().values():
                if not v.dtype.is_floating_point:
                    v.fill_(1234)
        broadcast_flat(flat, 0, 77)  # ONE collective: image + version + integer buffers
        ref = synth_state_dict(at.AlphaTriangleNet(mc, env).state_dict(), 42)
        ok = received_version(flat) == (77, flat.image_floats)
        for k, v in model.state_dict().items():
            if v.dtype.is_floating_point:
                ok = ok and torch.equal(v, ref[k])
            else:
                ok = ok and int(v) == 1234
        # rank 0 takes an optimiser step; the parameters are views of the image, so the next publication needs no packing
        if rank == 0:
            opt = torch.optim.SGD(model.parameters(), lr=0.1)
            model.train()
            pl, vl = model(torch.randn(4, 1, env.ROWS, env.COLS), torch.randn(4, mc.OTHER_NN_INPUT_FEATURES_DIM))
            (pl.square().mean() + vl.square().mean()).backward()
            opt.step()
            model.eval()
        before = flat.image.clone()
        broadcast_flat(flat, 0, 78)
        changed = not torch.equal(before, flat.image)
        ok = ok and (changed if rank != 0 else True) and received_version(flat)[0] == 78 and flat.aliased()
        gathered = [None] * world
        dist.all_gather_object(gathered, float(flat.image.double().sum()))
        ok = ok and gathered[0] == gathered[1]
        # each rank owns a disjoint shard of the games: nothing else is exchanged
        shard = list(range(rank, 64, world))
        dist.all_gather_object(gathered, shard)
        ok = ok and sorted(sum(gathered, [])) == list(range(64))
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_weight_broadcast_world2_gloo():
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}


def test_parameters_live_in_the_image():
    mc, env = at.ModelConfig(USE_TRANSFORMER=False), at.EnvConfig()
    a = at.AlphaTriangleNet(mc, env)
    want = synth_state_dict(a.state_dict(), 3)
    flat = _flat(a, mc, env)
    assert sum(n for _, _, n in flat.table) == 2742843  # 10,971,452 B state_dict minus ten int64 num_batches_tracked
    assert all(off % 4 == 0 for _, off, _ in flat.table)  # 16-byte-aligned views
    stamp0 = flat.stamp()
    a.load_state_dict(want)  # copies in place -> into the image
    assert flat.stamp() != stamp0 and flat.aliased()
    for key, off, numel in flat.table:
        assert torch.equal(flat.image[off : off + numel], want[key].reshape(-1)), key
    # an optimiser step writes the image too
    opt = torch.optim.Adam(a.parameters(), lr=1e-2)
    a.train()
    pl, vl = a(torch.randn(3, 1, env.ROWS, env.COLS), torch.randn(3, mc.OTHER_NN_INPUT_FEATURES_DIM))
    (pl.mean() + vl.mean()).backward()
    snapshot = flat.image.clone()
    opt.step()
    assert not torch.equal(snapshot, flat.image) and flat.aliased()
    # replacing storage is detected, and adopt() re-binds
    a.double().float()
    assert not flat.aliased()
    flat.adopt()
    assert flat.aliased()
