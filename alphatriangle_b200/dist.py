"""Weight distribution for sharded leaf evaluation.

The path shards by game: every rank evaluates its own states with a full weight
replica, and the only exchange is the trainer publishing new weights. The
reference does that with `ray.put(get_weights())` + `set_weights.remote(...)` on
every worker while the workers keep playing (alphatriangle/training/worker_manager.py:169-209,
triggered from training/loop.py:271-287). Here it is ONE collective over ONE
pre-registered buffer: every rank's parameters already live inside a flat fp32
image (flat.py), whose int64 tail carries the version (trainer global_step) and the
BatchNorm `num_batches_tracked` counters -- nothing is packed before the send or
unpacked after the receive, and no value comes back to the host in between. The
receiver folds / repacks the image on its GPU straight out of that buffer into the
engine's inactive weight set on a side stream, while evaluations on the main
stream keep reading the old set (at_engine_publish_weights). Backend: NCCL over
NVLink for CUDA tensors; the same code runs over gloo with CPU tensors, which is
how the host logic is tested without GPUs.
"""

from __future__ import annotations

import time

import torch
import torch.distributed as dist

from .flat import FlatWeights


def broadcast_flat(flat: FlatWeights, src: int, version: int, group=None) -> None:
    """All ranks call this. One broadcast of image + header; receivers' module tensors (views of the blob) and integer
    buffers hold the sender's values afterwards. Device-side only: no `.item()`, no host synchronisation."""
    if dist.get_rank(group) == src:
        flat.write_header(version)
    dist.broadcast(flat.blob, src=src, group=group)
    if dist.get_rank(group) != src:
        flat.apply_header()


def received_version(flat: FlatWeights) -> tuple[int, int]:
    """(version, image floats) as the sender wrote them; synchronises with the device (call it off the hot path)."""
    h = flat.header[:2].cpu()
    return int(h[0]), int(h[1])


class Publisher:
    """Per-process state of overlapped weight publication: a side stream for the collective and the fold / repack."""

    def __init__(self, nn, group=None):
        self.nn = nn
        self.group = group
        self.stream = torch.cuda.Stream(device=nn.device)
        self.done = torch.cuda.Event()

    def publish(self, src: int = 0, version: int | None = None) -> None:
        """Queue one publication and return. On the sender the side stream first waits for the work queued so far on the
        current stream (the optimiser step that produced the weights); receivers' evaluations are not waited for and keep
        reading their current set until the new one is ready."""
        nn = self.nn
        ver = nn._weights_version + 1 if version is None else int(version)
        if dist.get_rank(self.group) == src:
            self.stream.wait_stream(torch.cuda.current_stream(nn.device))
        with torch.cuda.stream(self.stream):
            broadcast_flat(nn.flat, src, ver, self.group)
            nn._weights_version = ver
            nn.engine.publish_image(nn.flat.image, ver)
            nn._stamp = nn.flat.stamp()
            self.done.record(self.stream)

    def wait(self) -> None:
        self.done.synchronize()


def broadcast_weights(nn, src: int = 0, version: int | None = None, group=None) -> float:
    """Publish rank `src`'s weights to every rank's engine on the current stream and wait. Returns this rank's elapsed
    milliseconds (collective + fold / repack, device idle otherwise)."""
    dev = nn.device
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    ver = nn._weights_version + 1 if version is None else int(version)
    broadcast_flat(nn.flat, src, ver, group)
    nn._weights_version = ver
    nn.engine.load_image(nn.flat.image, ver)
    nn._stamp = nn.flat.stamp()
    torch.cuda.synchronize(dev)
    return (time.perf_counter() - t0) * 1e3
