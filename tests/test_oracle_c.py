"""oracle/features_ref.c (plain C, float64 arithmetic) against the vectors the reference produced (tests/golden/) and
against the numpy restatement on fresh states: a third independent statement of the bit-exact feature contract."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest
from golden_cases import CASES, case_configs

import alphatriangle_b200 as at
from alphatriangle_b200.packing import pack_states_py
from alphatriangle_b200.synth import edge_states, synth_states
from oracle import features_ref

LIB = Path(__file__).resolve().parents[1] / "oracle" / "liboracle_features.so"


@pytest.fixture(scope="module")
def lib():
    if not LIB.exists():
        pytest.fail(f"{LIB} not built: run `python -c 'import __graft_entry__ as g; g.build()'` (make -C oracle)")
    lib = C.CDLL(str(LIB))
    lib.at_oracle_features.restype = C.c_int
    lib.at_oracle_features.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return lib


def run(lib, packed, rows, cols, slots, channels, dim):
    packed = np.ascontiguousarray(packed)
    n = packed.shape[0]
    grid = np.empty((n, channels, rows, cols), dtype=np.float32)
    other = np.empty((n, dim), dtype=np.float32)
    assert lib.at_oracle_features(packed.ctypes.data, n, rows, cols, slots, channels, dim, grid.ctypes.data, other.ctypes.data) == 0
    return grid, other


@pytest.mark.parametrize("case", sorted(CASES))
def test_c_oracle_reproduces_the_reference_vectors(lib, case, golden_dir):
    mc_kwargs, env_kwargs, _seed = case_configs(case)
    mc, env = at.ModelConfig(**mc_kwargs), at.EnvConfig(**env_kwargs)
    g = np.load(golden_dir / f"{case}.npz")
    grid, other = run(lib, g["packed"], env.ROWS, env.COLS, env.NUM_SHAPE_SLOTS, mc.GRID_INPUT_CHANNELS, mc.OTHER_NN_INPUT_FEATURES_DIM)
    assert np.array_equal(grid.view(np.uint32), g["grid"].view(np.uint32))
    assert np.array_equal(other.view(np.uint32), g["other"].view(np.uint32))


def test_c_oracle_matches_numpy_oracle_on_fresh_and_edge_states(lib):
    env, mc = at.EnvConfig(), at.ModelConfig()
    states = synth_states(500, env, seed=99) + edge_states(env)
    grid_np, other_np = features_ref.extract_batch(states, mc)
    grid, other = run(lib, pack_states_py(states, env), env.ROWS, env.COLS, env.NUM_SHAPE_SLOTS, mc.GRID_INPUT_CHANNELS,
                      mc.OTHER_NN_INPUT_FEATURES_DIM)
    assert np.array_equal(grid.view(np.uint32), grid_np.view(np.uint32))
    assert np.array_equal(other.view(np.uint32), other_np.view(np.uint32))
    # truncation and zero padding of the feature vector (extractor.py:127-139)
    for dim in (5, 24, 40):
        mc2 = at.ModelConfig(OTHER_NN_INPUT_FEATURES_DIM=dim)
        _g, o_np = features_ref.extract_batch(states[:50], mc2)
        _g2, o_c = run(lib, pack_states_py(states[:50], env), env.ROWS, env.COLS, env.NUM_SHAPE_SLOTS, 1, dim)
        assert np.array_equal(o_c.view(np.uint32), o_np.view(np.uint32))
    assert lib.at_oracle_features(None, 0, 20, 20, 3, 1, 30, None, None) == -1
