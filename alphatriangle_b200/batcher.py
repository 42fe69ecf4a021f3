"""Cross-game leaf batching.

The reference hands one NeuralNetwork to one trimcts search, which calls evaluate_batch with at most
`mcts_batch_size` leaves at a time (32 by default: alphatriangle/config/mcts_config.py:14,58-62, call site
alphatriangle/rl/self_play/worker.py:273-279). On a GPU whose evaluate step is flat up to thousands of states that
wastes almost all of it. LeafBatcher sits between many concurrent searches (one thread each) and one NeuralNetwork:
every search gets a `client()` that looks like the network (evaluate_state / evaluate_batch, the
trimcts.AlphaZeroNetworkInterface surface), requests from all clients are coalesced into one evaluate_batch call of up
to `max_batch` states, and each caller gets back exactly the results for its own states, in order.

A batch is dispatched when (a) `max_batch` states are waiting, or (b) every open client is blocked in a request -- no
one is left who could add to the batch -- or (c) the oldest request has waited `max_wait_s`. Errors of a coalesced call
are re-raised in every caller that had states in it.
"""

from __future__ import annotations

import threading
import time
from typing import Any, Sequence

__all__ = ["LeafBatcher", "BatcherClient", "BatcherClosed"]


class BatcherClosed(RuntimeError):
    """Raised in callers that are waiting (or arrive) after LeafBatcher.close()."""


class _Request:
    __slots__ = ("states", "results", "error", "done", "t0")

    def __init__(self, states):
        self.states = states
        self.results = None
        self.error: BaseException | None = None
        self.done = False
        self.t0 = time.monotonic()


class BatcherClient:
    """What one search holds instead of the NeuralNetwork: same evaluate_state / evaluate_batch contract."""

    def __init__(self, owner: "LeafBatcher"):
        self._owner = owner
        self._open = True

    def evaluate_batch(self, states: Sequence[Any]):
        if not states:
            return []
        return self._owner._submit(list(states))

    def evaluate_state(self, state: Any):
        return self._owner._submit([state])[0]

    def close(self) -> None:
        if self._open:
            self._open = False
            self._owner._client_closed()

    def __enter__(self) -> "BatcherClient":
        return self

    def __exit__(self, *exc) -> None:
        self.close()


class LeafBatcher:
    def __init__(self, network, max_batch: int = 4096, max_wait_s: float = 0.002):
        if max_batch <= 0:
            raise ValueError("max_batch must be positive")
        self._nn = network
        self.max_batch = int(max_batch)
        self.max_wait_s = float(max_wait_s)
        self._cv = threading.Condition()
        self._pending: list[_Request] = []
        self._pending_states = 0
        self._clients = 0  # open clients
        self._closed = False
        # statistics (read them after close(), or under no concurrency)
        self.calls = 0
        self.states = 0
        self.requests = 0
        self._thread = threading.Thread(target=self._run, name="LeafBatcher", daemon=True)
        self._thread.start()

    # ------------------------------------------------------------------ client side
    def client(self) -> BatcherClient:
        with self._cv:
            if self._closed:
                raise BatcherClosed("LeafBatcher is closed")
            self._clients += 1
        return BatcherClient(self)

    def _client_closed(self) -> None:
        with self._cv:
            self._clients -= 1
            self._cv.notify_all()  # the "everyone is waiting" rule may hold now

    def _submit(self, states: list):
        req = _Request(states)
        with self._cv:
            if self._closed:
                raise BatcherClosed("LeafBatcher is closed")
            self._pending.append(req)
            self._pending_states += len(states)
            self.requests += 1
            self._cv.notify_all()
            while not req.done:
                self._cv.wait()
        if req.error is not None:
            raise req.error
        return req.results

    # ------------------------------------------------------------------ dispatcher
    def _ready(self) -> bool:
        if not self._pending:
            return False
        if self._pending_states >= self.max_batch:
            return True
        if len(self._pending) >= max(self._clients, 1):  # nobody left who could add to this batch
            return True
        return time.monotonic() - self._pending[0].t0 >= self.max_wait_s

    def _take(self) -> list[_Request]:
        """Whole requests, oldest first, up to max_batch states (a single larger request goes alone)."""
        taken: list[_Request] = []
        n = 0
        while self._pending and (not taken or n + len(self._pending[0].states) <= self.max_batch):
            r = self._pending.pop(0)
            taken.append(r)
            n += len(r.states)
        self._pending_states -= n
        return taken

    def _run(self) -> None:
        while True:
            with self._cv:
                while not self._closed and not self._ready():
                    if self._pending:
                        left = self.max_wait_s - (time.monotonic() - self._pending[0].t0)
                        self._cv.wait(timeout=max(left, 0.0) + 1e-4)
                    else:
                        self._cv.wait()
                if self._closed:
                    for r in self._pending:
                        r.error = BatcherClosed("LeafBatcher was closed while the request was waiting")
                        r.done = True
                    self._pending.clear()
                    self._pending_states = 0
                    self._cv.notify_all()
                    return
                batch = self._take()
            flat = [s for r in batch for s in r.states]
            try:
                out = self._nn.evaluate_batch(flat)
                if len(out) != len(flat):
                    raise RuntimeError(f"network returned {len(out)} results for {len(flat)} states")
                at = 0
                for r in batch:
                    r.results = out[at : at + len(r.states)]
                    at += len(r.states)
            except BaseException as e:  # every caller of this batch sees the failure
                for r in batch:
                    r.error = e
            with self._cv:
                self.calls += 1
                self.states += len(flat)
                for r in batch:
                    r.done = True
                self._cv.notify_all()

    def close(self) -> None:
        with self._cv:
            self._closed = True
            self._cv.notify_all()
        self._thread.join(timeout=10.0)

    def __enter__(self) -> "LeafBatcher":
        return self

    def __exit__(self, *exc) -> None:
        self.close()
