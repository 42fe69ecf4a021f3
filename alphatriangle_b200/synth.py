"""Seeded synthetic game states.

`trianglengin` is a compiled third-party package that is not available where this
repository is built and measured, so tests, the benchmark and the CPU baseline
all draw states from this generator. `SynthState` exposes exactly the accessor
surface the evaluation path touches (extractor.py:25-31,58-66,109,114 and
network.py:201,289) and nothing else.
"""

from __future__ import annotations

import numpy as np

from .config import playable_mask


class SynthShape:
    __slots__ = ("triangles", "color")

    def __init__(self, triangles):
        self.triangles = list(triangles)
        self.color = (0, 0, 0)

    def bbox(self):
        if not self.triangles:
            return (0, 0, 0, 0)
        rs = [t[0] for t in self.triangles]
        cs = [t[1] for t in self.triangles]
        return (min(rs), min(cs), max(rs), max(cs))


class SynthState:
    def __init__(self, env_config, occupied, death, shapes, score, step, valid=None):
        self.env_config = env_config
        self._occupied = occupied
        self._death = death
        self._shapes = shapes
        self._score = score
        self.current_step = step
        self._valid = valid

    def get_grid_data_np(self):
        return {"occupied": self._occupied, "death": self._death, "color_id": np.zeros(self._occupied.shape, dtype=np.int8)}

    def get_shapes(self):
        return list(self._shapes)

    def game_score(self):
        return self._score

    def valid_actions(self):
        if self._valid is not None:
            return list(self._valid)
        return list(range(int(self.env_config.NUM_SHAPE_SLOTS * self.env_config.ROWS * self.env_config.COLS)))

    def is_over(self):
        return False

    def copy(self):
        return SynthState(self.env_config, self._occupied.copy(), self._death.copy(), list(self._shapes), self._score,
                          self.current_step, self._valid)

    def __repr__(self):
        return f"SynthState(step={self.current_step}, score={self._score:.3f})"


def _polyiamond(rng: np.random.Generator, n: int) -> SynthShape:
    """Connected set of n triangles grown from (0,0); up-triangles have (r+c) even and a neighbour below."""
    cells = [(0, 0)]
    while len(cells) < n:
        r, c = cells[int(rng.integers(len(cells)))]
        up = (r + c) % 2 == 0
        cand = [(r, c - 1), (r, c + 1), (r + 1, c) if up else (r - 1, c)]
        nxt = cand[int(rng.integers(3))]
        if nxt not in cells:
            cells.append(nxt)
    min_r = min(r for r, _ in cells)
    min_c = min(c for _, c in cells)
    return SynthShape([(r - min_r, c - min_c, (r + c) % 2 == 0) for r, c in cells])


def synth_states(n: int, env_config, seed: int = 1234) -> list[SynthState]:
    rng = np.random.default_rng(seed)
    rows, cols, slots = int(env_config.ROWS), int(env_config.COLS), int(env_config.NUM_SHAPE_SLOTS)
    death = ~np.array(playable_mask(env_config), dtype=bool)
    out = []
    for _ in range(n):
        p = rng.uniform(0.0, 0.9)
        occupied = (rng.random((rows, cols)) < p) & ~death
        shapes = []
        for _s in range(slots):
            if rng.random() < 0.2:
                shapes.append(None)
            else:
                shapes.append(_polyiamond(rng, int(rng.integers(1, 6))))
        score = float(rng.normal(20.0, 30.0))
        step = int(rng.integers(0, 1200))
        out.append(SynthState(env_config, occupied, death.copy(), shapes, score, step))
    return out


def edge_states(env_config) -> list[SynthState]:
    """Hand-built corner cases: empty / full boards, empty slots, clipped score and step, NaN / inf score,
    occupied bits on death cells, a present-but-empty shape, shapes wider than the board."""
    rows, cols, slots = int(env_config.ROWS), int(env_config.COLS), int(env_config.NUM_SHAPE_SLOTS)
    death = ~np.array(playable_mask(env_config), dtype=bool)
    empty = np.zeros((rows, cols), dtype=bool)
    full = ~death
    everything = np.ones((rows, cols), dtype=bool)
    tri1 = SynthShape([(0, 0, True)])
    big = SynthShape([(0, 0, True), (0, 1, False), (0, 2, True), (1, 1, False), (1, 2, True)])
    wide = SynthShape([(0, c, c % 2 == 0) for c in range(cols + 3)])
    tall = SynthShape([(r, 0, r % 2 == 0) for r in range(rows + 2)])
    hollow = SynthShape([])
    neg = SynthShape([(-2, -3, True), (-1, -3, False)])
    top_row = empty.copy()
    top_row[rows - 1, :] = True
    top_row &= ~death
    checker = (np.add.outer(np.arange(rows), np.arange(cols)) % 2 == 0) & ~death
    col0 = empty.copy()
    col0[:, cols // 2] = True
    col0 &= ~death

    def mk(occ, shapes, score, step):
        shapes = (list(shapes) + [None] * slots)[:slots]
        return SynthState(env_config, occ.copy(), death.copy(), shapes, score, step)

    return [
        mk(empty, [None] * slots, 0.0, 0),
        mk(full, [tri1] * slots, 499.999, 999),
        mk(full, [big, wide, tall], 500.0, 1000),
        mk(everything, [hollow, neg, big], -500.1, 1001),
        mk(top_row, [wide, None, tri1], 1e308, 10**12),
        mk(checker, [tall, hollow, None], -1e308, -5),
        mk(col0, [neg, neg, neg], float("nan"), 1),
        mk(checker, [big, None, None], float("inf"), 500),
        mk(top_row, [None, big, None], float("-inf"), 7),
        mk(col0, [tri1, big, wide], -0.0, 123),
        mk(empty, [big] * slots, 1e-320, 2**31 - 1),
        mk(full, [None] * slots, 33.333333333333336, -(2**31)),
    ]


def synth_state_dict(module_state_dict: dict, seed: int = 0, sharpness: float = 0.0) -> dict:
    """Deterministic weights for a reference-keyed state_dict, independent of torch's init order.

    Every float tensor is drawn from numpy's PCG64 in sorted-key order: weights ~ U(-a, a) with
    a = 1/sqrt(fan_in); BatchNorm running_mean ~ N(0, 0.1), running_var ~ U(0.5, 1.5), affine
    weight ~ U(0.8, 1.2), bias ~ N(0, 0.05) so that folding is really exercised; LayerNorm likewise.
    `pos_encoder.pe` and integer buffers are left as they are.

    sharpness = 0 gives the small weights above: activations shrink with depth and the policy comes out
    uniform to three digits (entropy ln 360 - 0.003), which says little about a trained network.
    sharpness > 0 gives trained-like statistics instead: matrices ~ U(-a, a) with a = sqrt(6 / fan_in)
    (variance-preserving through ReLU, so activations stay O(1) through the depth), and the last Linear of
    the policy and value heads scaled by `sharpness`, which sets the spread of the logits: 2 gives policies
    of 2-3 nats with a clear best move, 8 near-deterministic ones; expected values spread over the support.
    """
    import torch

    rng = np.random.default_rng(seed)
    sharp = float(sharpness) > 0.0
    last_linear = {}
    if sharp:
        for head in ("policy_head", "value_head"):
            mats = sorted((int(k.split(".")[1]), k) for k, t in module_state_dict.items()
                          if k.startswith(head + ".") and k.endswith(".weight") and t.dim() == 2)
            if mats:
                last_linear[mats[-1][1]] = float(sharpness)
    out = {}
    for key in sorted(module_state_dict):
        t = module_state_dict[key]
        if not t.dtype.is_floating_point or key == "pos_encoder.pe":
            out[key] = t.clone()
            continue
        shape = tuple(t.shape)
        leaf = key.rsplit(".", 1)[-1]
        if leaf == "running_mean":
            a = rng.normal(0.0, 0.1, shape)
        elif leaf == "running_var":
            a = rng.uniform(0.5, 1.5, shape)
        elif len(shape) == 1 and leaf == "weight":
            a = rng.uniform(0.8, 1.2, shape)
        elif len(shape) == 1:
            a = rng.normal(0.0, 0.05, shape)
        else:
            fan_in = int(np.prod(shape[1:]))
            bound = np.sqrt(6.0 / fan_in) if sharp else 1.0 / np.sqrt(fan_in)
            a = rng.uniform(-bound, bound, shape) * last_linear.get(key, 1.0)
        out[key] = torch.from_numpy(a.astype(np.float32)).to(t.device)
    return out
