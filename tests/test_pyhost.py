"""C host loops (csrc/at_pyhost.c) against their Python twins: SURVEY.md 8f rows 1-2."""
import numpy as np
import pytest

import alphatriangle_b200 as at
from alphatriangle_b200 import packing
from alphatriangle_b200.synth import edge_states, synth_states


def test_c_packer_matches_the_python_twin_bit_for_bit():
    env = at.EnvConfig()
    states = synth_states(300, env, seed=31) + edge_states(env)
    a = packing.pack_states(states, env)
    b = packing.pack_states_py(states, env)
    assert a.dtype == np.uint8 and a.shape == (len(states), packing.RECORD_BYTES)
    assert np.array_equal(a, b)


def test_c_packer_other_board_sizes_and_into_out():
    for rows, cols, slots in ((3, 4, 1), (8, 16, 4), (5, 7, 2)):
        env = at.EnvConfig(ROWS=rows, COLS=cols, NUM_SHAPE_SLOTS=slots, PLAYABLE_RANGE_PER_ROW=[(0, cols)] * rows)
        states = synth_states(40, env, seed=rows * 100 + cols)
        out = np.full((len(states), packing.RECORD_BYTES), 0xAB, dtype=np.uint8)
        got = packing.pack_states(states, env, out)
        assert got.base is not None and np.shares_memory(got, out)
        assert np.array_equal(out, packing.pack_states_py(states, env))


class _ListGridState:
    """Grid arrays of a wider dtype / Fortran order: the C loop reads strided byte arrays itself and hands anything else back."""

    def __init__(self, s, wide):
        self._s, self._wide = s, wide
        self.current_step = s.current_step

    def get_grid_data_np(self):
        g = self._s.get_grid_data_np()
        if self._wide:
            return {"occupied": g["occupied"].astype(np.int64), "death": g["death"].astype(np.float32)}
        return {"occupied": np.asfortranarray(g["occupied"]), "death": np.asfortranarray(g["death"])}

    def get_shapes(self):
        return self._s.get_shapes()

    def game_score(self):
        return self._s.game_score()


def test_c_packer_hands_unusual_arrays_to_the_python_packer():
    env = at.EnvConfig()
    base = synth_states(12, env, seed=5)
    ref = packing.pack_states_py(base, env)
    for wide in (False, True):
        mixed = list(base[:5]) + [_ListGridState(s, wide) for s in base[5:]]
        assert np.array_equal(packing.pack_states(mixed, env), ref)


def test_c_packer_errors_match():
    env = at.EnvConfig()
    s = synth_states(1, env, seed=1)[0]

    class Bad:
        current_step = 0

        def get_grid_data_np(self):
            g = s.get_grid_data_np()
            return {"occupied": g["occupied"][:, :-1], "death": g["death"][:, :-1]}

        def get_shapes(self):
            return s.get_shapes()

        def game_score(self):
            return 0.0

    with pytest.raises(ValueError, match="expected"):
        packing.pack_states([Bad()], env)

    class TooMany(Bad):
        def get_grid_data_np(self):
            return s.get_grid_data_np()

        def get_shapes(self):
            return [None] * (env.NUM_SHAPE_SLOTS + 1)

    with pytest.raises(ValueError, match="shape slots"):
        packing.pack_states([TooMany()], env)

    class Raises(TooMany):
        def get_shapes(self):
            raise RuntimeError("engine gone")

    with pytest.raises(RuntimeError, match="engine gone"):
        packing.pack_states([Raises()], env)


def test_c_packer_step_saturates_and_score_converts():
    env = at.EnvConfig()
    s = synth_states(1, env, seed=2)[0]
    s.current_step = 2**40
    s._score = 7  # an int score, as float(state.game_score()) accepts
    rec = packing.pack_states([s], env).reshape(-1).view(packing.PACKED_DTYPE)[0]
    assert rec["step"] == 2**31 - 1 and rec["score"] == 7.0
    s.current_step = -(2**40)
    rec = packing.pack_states([s], env).reshape(-1).view(packing.PACKED_DTYPE)[0]
    assert rec["step"] == -(2**31)


def test_result_builder_matches_dict_zip():
    host = packing._host()
    rng = np.random.default_rng(3)
    n, a = 37, 360
    probs = rng.random((n, a), dtype=np.float32)
    probs /= probs.sum(axis=1, keepdims=True)
    probs[3, 11] += probs[3, 10]
    probs[3, 10] = -1e-9  # clipped to zero, the row still sums to one
    probs[7] *= 0.9  # off by more than the gate: reported for the exact rule
    probs[9, 0] = np.nan
    values = rng.standard_normal(n).astype(np.float32)
    keys = tuple(range(a))
    results, recheck = host.build_results(probs, values, n, a, keys)
    assert recheck == [7, 9]
    assert len(results) == n
    for i in range(n):
        d, v = results[i]
        assert type(d) is dict and type(v) is float and list(d.keys()) == list(keys)
        assert v == float(values[i])
        if i != 9:
            assert list(d.values()) == np.maximum(probs[i], 0).tolist()
    assert results[3][0][10] == 0.0 and all(type(x) is float for x in results[0][0].values())
    assert results[9][0][0] == 0.0


def test_result_builder_rejects_wrong_buffers():
    host = packing._host()
    keys = tuple(range(4))
    with pytest.raises(ValueError):
        host.build_results(np.zeros((2, 4), np.float64), np.zeros(2, np.float32), 2, 4, keys)
    with pytest.raises(ValueError):
        host.build_results(np.zeros((2, 4), np.float32), np.zeros(3, np.float32), 2, 4, keys)
    with pytest.raises(ValueError):
        host.build_results(np.zeros((2, 4), np.float32), np.zeros(2, np.float32), 2, 4, tuple(range(5)))
    assert host.build_results(np.zeros((0, 4), np.float32), np.zeros(0, np.float32), 0, 4, keys) == ([], [])
