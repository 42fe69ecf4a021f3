"""Weight distribution for sharded leaf evaluation.

The path shards by game: every rank evaluates its own states with a full weight
replica, and the only exchange is the trainer publishing new weights. The
reference does that with `ray.put(get_weights())` + `set_weights.remote(...)` on
every worker (alphatriangle/training/worker_manager.py:169-209, triggered from
training/loop.py:271-287). Here it is ONE broadcast of the flat fp32 weight
image (order = at_model_describe) plus a small int64 header (version =
trainer global_step, and the BatchNorm `num_batches_tracked` counters, which are
the only non-float entries of the state_dict). The receiver folds / repacks the
image on its GPU straight out of the receive buffer (at_engine_load_weights)
into the engine's inactive weight set, so in-flight batches keep reading the old
set. Backend: NCCL over NVLink for CUDA tensors; the same code runs over gloo
with CPU tensors, which is how the host logic is tested without GPUs.
"""

from __future__ import annotations

import time

import torch
import torch.distributed as dist


def float_table(state_dict, table):
    """[(key, offset, numel)] restricted / checked against a state_dict."""
    for key, _off, numel in table:
        if key not in state_dict or state_dict[key].numel() != numel:
            raise KeyError(f"state_dict does not match the weight table at {key!r}")
    return table


def int_keys(state_dict) -> list[str]:
    return [k for k, v in state_dict.items() if not v.dtype.is_floating_point]


def pack_image(state_dict, table, device) -> torch.Tensor:
    parts = [state_dict[k].detach().to(device=device, dtype=torch.float32).reshape(-1) for k, _o, _n in float_table(state_dict, table)]
    return torch.cat(parts) if parts else torch.empty(0, dtype=torch.float32, device=device)


def unpack_image(image: torch.Tensor, table, state_dict) -> None:
    """Copy the image back into the tensors of `state_dict` in place (keeps get_weights() truthful on receivers)."""
    with torch.no_grad():
        for key, off, numel in table:
            t = state_dict[key]
            t.copy_(image[off : off + numel].reshape(t.shape).to(device=t.device, dtype=t.dtype))


def broadcast_state(state_dict, table, src: int, version: int, device, group=None):
    """All ranks call this. Returns (image on `device`, version). Non-src ranks get their int buffers overwritten."""
    rank = dist.get_rank(group)
    ikeys = int_keys(state_dict)
    total = sum(n for _k, _o, n in table)
    header = torch.zeros(2 + len(ikeys), dtype=torch.int64, device=device)
    if rank == src:
        header[0] = version
        header[1] = total
        for i, k in enumerate(ikeys):
            header[2 + i] = state_dict[k].reshape(-1)[0].to(torch.int64)
        image = pack_image(state_dict, table, device)
    else:
        image = torch.empty(total, dtype=torch.float32, device=device)
    dist.broadcast(header, src=src, group=group)
    if int(header[1].item()) != total:
        raise RuntimeError(f"weight image size mismatch: sender has {int(header[1].item())} floats, this rank expects {total}")
    dist.broadcast(image, src=src, group=group)
    if rank != src:
        with torch.no_grad():
            for i, k in enumerate(ikeys):
                state_dict[k].fill_(int(header[2 + i].item()))
    return image, int(header[0].item())


def broadcast_weights(nn, src: int = 0, version: int | None = None, group=None) -> float:
    """Publish rank `src`'s weights to every rank's engine. Returns this rank's elapsed milliseconds."""
    dev = nn.device
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    sd = nn._orig_model.state_dict()
    ver = nn._weights_version + 1 if version is None else version
    image, ver = broadcast_state(sd, nn.engine.table, src, ver, dev, group)
    if dist.get_rank(group) != src:
        unpack_image(image, nn.engine.table, sd)
        nn._stamp = nn._module_stamp()
    nn._weights_version = ver
    nn.engine.load_image(image, ver)
    nn._stamp = nn._module_stamp()
    torch.cuda.synchronize(dev)
    return (time.perf_counter() - t0) * 1e3
