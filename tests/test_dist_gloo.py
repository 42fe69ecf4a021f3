"""world_size-2 gloo run of the weight-publication host logic (CPU tensors, no GPU)."""

import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import alphatriangle_b200 as at
from alphatriangle_b200 import _lib
from alphatriangle_b200.dist import broadcast_state, pack_image, unpack_image
from alphatriangle_b200.synth import synth_state_dict


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mc = at.ModelConfig(CONV_FILTERS=[8, 16], CONV_KERNEL_SIZES=[3, 3], CONV_STRIDES=[1, 1], CONV_PADDING=[1, 1],
                            NUM_RESIDUAL_BLOCKS=1, RESIDUAL_BLOCK_FILTERS=16, TRANSFORMER_DIM=16, TRANSFORMER_HEADS=2,
                            TRANSFORMER_LAYERS=1, TRANSFORMER_FC_DIM=32)
        env = at.EnvConfig()
        table = _lib.weight_table(mc, env)
        torch.manual_seed(100 + rank)  # ranks start with different weights
        model = at.AlphaTriangleNet(mc, env)
        sd = model.state_dict()
        if rank == 0:
            model.load_state_dict(synth_state_dict(sd, 42))
            for k, v in model.state_dict().items():
                if not v.dtype.is_floating_point:
                    v.fill_(1234)
        sd = model.state_dict()
        image, version = broadcast_state(sd, table, 0, 77, torch.device("cpu"))
        if rank != 0:
            unpack_image(image, table, sd)
        ref = synth_state_dict(at.AlphaTriangleNet(mc, env).state_dict(), 42)
        ok = version == 77
        for k, v in model.state_dict().items():
            if v.dtype.is_floating_point:
                ok = ok and torch.equal(v, ref[k])
            else:
                ok = ok and int(v) == 1234
        ok = ok and torch.equal(image, pack_image(ref, table, torch.device("cpu")))
        # each rank owns a disjoint shard of the games: nothing else is exchanged
        shard = list(range(rank, 64, world))
        gathered = [None] * world
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


def test_pack_unpack_round_trip():
    mc, env = at.ModelConfig(USE_TRANSFORMER=False), at.EnvConfig()
    table = _lib.weight_table(mc, env)
    a = at.AlphaTriangleNet(mc, env)
    b = at.AlphaTriangleNet(mc, env)
    a.load_state_dict(synth_state_dict(a.state_dict(), 3))
    image = pack_image(a.state_dict(), table, torch.device("cpu"))
    assert image.numel() == sum(n for _, _, n in table) == 2742843  # 10,971,452 B state_dict minus ten int64 num_batches_tracked
    unpack_image(image, table, b.state_dict())
    for k, v in a.state_dict().items():
        if v.dtype.is_floating_point:
            assert torch.equal(v, b.state_dict()[k])
