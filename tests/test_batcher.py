"""Cross-game leaf batching (alphatriangle_b200/batcher.py): host logic only, a stand-in network."""
import threading
import time

import pytest

from alphatriangle_b200.batcher import BatcherClosed, LeafBatcher


class FakeNet:
    """evaluate_batch echoes each state's id; records the size of every call."""

    def __init__(self, delay=0.0, fail_on=None):
        self.sizes = []
        self.delay = delay
        self.fail_on = fail_on

    def evaluate_batch(self, states):
        self.sizes.append(len(states))
        if self.delay:
            time.sleep(self.delay)
        if self.fail_on is not None and any(s == self.fail_on for s in states):
            raise ValueError(f"poisoned state {self.fail_on}")
        return [({0: float(s)}, float(s) * 2) for s in states]


def _game(client, gid, n_requests, leaves, out, errors):
    try:
        with client:
            for r in range(n_requests):
                states = [gid * 1_000_000 + r * 1000 + k for k in range(leaves)]
                res = client.evaluate_batch(states)
                assert [int(d[0]) for d, _v in res] == states
                assert [v for _d, v in res] == [float(s) * 2 for s in states]
                out.append(len(res))
    except BaseException as e:  # noqa: BLE001
        errors.append(e)


def test_requests_of_many_games_are_coalesced_and_routed_back():
    net = FakeNet(delay=0.0005)
    games, reqs, leaves = 24, 20, 32
    with LeafBatcher(net, max_batch=4096, max_wait_s=0.05) as b:
        clients = [b.client() for _ in range(games)]
        out, errors = [], []
        threads = [threading.Thread(target=_game, args=(c, g, reqs, leaves, out, errors)) for g, c in enumerate(clients)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=60)
        assert not errors, errors
        assert sum(out) == games * reqs * leaves
        assert b.states == games * reqs * leaves and b.requests == games * reqs
        # every round all 24 games block, so a call carries (about) 24 requests, far fewer calls than requests
        assert b.calls <= b.requests // 8
        assert max(net.sizes) <= 4096 and max(net.sizes) >= 12 * leaves


def test_max_batch_splits_and_single_state_calls():
    net = FakeNet()
    with LeafBatcher(net, max_batch=64, max_wait_s=0.01) as b:
        clients = [b.client() for _ in range(5)]
        out, errors = [], []
        threads = [threading.Thread(target=_game, args=(c, g, 6, 32, out, errors)) for g, c in enumerate(clients)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=60)
        assert not errors, errors
        assert max(net.sizes) <= 64
        with b.client() as c:
            assert c.evaluate_state(7) == ({0: 7.0}, 14.0)
            assert c.evaluate_batch([]) == []
            big = list(range(200))  # larger than max_batch: goes alone, whole
            assert [int(d[0]) for d, _ in c.evaluate_batch(big)] == big


def test_timeout_dispatches_when_some_games_are_busy_elsewhere():
    net = FakeNet()
    with LeafBatcher(net, max_batch=4096, max_wait_s=0.02) as b:
        idle = b.client()  # open, never asks: rule (b) cannot fire
        with b.client() as c:
            t0 = time.monotonic()
            assert c.evaluate_state(3)[1] == 6.0
            assert 0.015 <= time.monotonic() - t0 < 1.0
        idle.close()


def test_error_reaches_every_caller_of_the_batch_and_later_batches_work():
    net = FakeNet(fail_on=13)
    with LeafBatcher(net, max_batch=4096, max_wait_s=0.05) as b:
        c1, c2 = b.client(), b.client()
        errs = []

        def ask(c, states):
            try:
                c.evaluate_batch(states)
            except ValueError as e:
                errs.append(str(e))

        t1 = threading.Thread(target=ask, args=(c1, [1, 2]))
        t2 = threading.Thread(target=ask, args=(c2, [13]))
        t1.start(); t2.start(); t1.join(10); t2.join(10)
        assert len(errs) == 2 and all("poisoned" in e for e in errs)
        net.fail_on = None
        c2.close()
        assert c1.evaluate_state(5) == ({0: 5.0}, 10.0)
        c1.close()


def test_close_releases_waiters():
    class Stuck:
        def __init__(self):
            self.gate = threading.Event()

        def evaluate_batch(self, states):
            self.gate.wait(5)
            return [({}, 0.0) for _ in states]

    net = Stuck()
    b = LeafBatcher(net, max_batch=8, max_wait_s=0.001)
    c1, c2 = b.client(), b.client()
    got = []

    def ask(c, s):
        try:
            got.append(c.evaluate_batch(s))
        except BatcherClosed as e:
            got.append(e)

    t1 = threading.Thread(target=ask, args=(c1, [1] * 8))  # taken by the dispatcher, stuck in the network
    t1.start()
    time.sleep(0.05)
    t2 = threading.Thread(target=ask, args=(c2, [2]))  # still queued when close() comes
    t2.start()
    time.sleep(0.05)
    closer = threading.Thread(target=b.close)
    closer.start()
    time.sleep(0.05)
    net.gate.set()
    t1.join(10); t2.join(10); closer.join(10)
    assert len(got) == 2 and any(isinstance(g, BatcherClosed) for g in got) and any(isinstance(g, list) for g in got)
    with pytest.raises(BatcherClosed):
        b.client()
