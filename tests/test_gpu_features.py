"""Feature kernel parity through the C ABI: bit-exact against the reference's recorded
outputs (golden) and against the oracle on fresh seeded states, plus size-independent
properties at BASELINE batch sizes."""

import numpy as np
import pytest
import torch
from golden_cases import CASES
from gpu_util import bits, load_case

import alphatriangle_b200 as at
from alphatriangle_b200.features import extract_features_batch, extract_state_features
from alphatriangle_b200.packing import pack_states
from alphatriangle_b200.synth import edge_states, synth_states
from oracle import features_ref

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", CASES)
def test_features_bit_exact_vs_reference_golden(name, golden_dir):
    mc, env, _seed, g = load_case(name, golden_dir)
    from alphatriangle_b200.features import _run

    grid, other = _run(g["packed"], env, mc.GRID_INPUT_CHANNELS, mc.OTHER_NN_INPUT_FEATURES_DIM, torch.device("cuda:0"))
    assert np.array_equal(bits(grid), bits(g["grid"]))
    assert np.array_equal(bits(other), bits(g["other"]))


def test_features_bit_exact_vs_oracle_2048_states():
    env, mc = at.EnvConfig(), at.ModelConfig()
    states = synth_states(2048, env, seed=99) + edge_states(env)
    feats = extract_features_batch(states, mc)
    grid_o, other_o = features_ref.extract_batch(states, mc)
    assert np.array_equal(bits(np.stack([f["grid"] for f in feats])), bits(grid_o))
    assert np.array_equal(bits(np.stack([f["other_features"] for f in feats])), bits(other_o))


def test_single_state_api_and_empty_batch():
    env, mc = at.EnvConfig(), at.ModelConfig()
    s = synth_states(1, env, seed=5)[0]
    f = extract_state_features(s, mc)
    o = features_ref.extract_state_features(s, mc)
    assert f["grid"].shape == (1, 8, 15) and f["grid"].dtype == np.float32
    assert f["other_features"].shape == (30,) and f["other_features"].dtype == np.float32
    assert np.array_equal(bits(f["grid"]), bits(o["grid"])) and np.array_equal(bits(f["other_features"]), bits(o["other_features"]))
    assert extract_features_batch([], mc) == []


@pytest.mark.parametrize("batch", [4096, 65536])
def test_features_full_size_properties(batch):
    """At BASELINE sizes: tiling a small seeded batch must reproduce the small batch's rows exactly
    (every block / tail path gives the same answer), and the value ranges the domain guarantees hold."""
    env, mc = at.EnvConfig(), at.ModelConfig()
    from alphatriangle_b200.features import _run

    base_states = synth_states(509, env, seed=3) + edge_states(env)  # 521 rows: not a multiple of the block's 16 states
    base = pack_states(base_states, env)
    reps = -(-batch // base.shape[0])
    packed = np.tile(base, (reps, 1))[:batch]
    grid, other = _run(packed, env, 1, 30, torch.device("cuda:0"))
    g0, o0 = features_ref.extract_batch(base_states, mc)
    idx = np.arange(batch) % base.shape[0]
    assert np.array_equal(bits(grid), bits(g0)[idx])
    assert np.array_equal(bits(other), bits(o0)[idx])
    assert set(np.unique(grid)) <= {-1.0, 0.0, 1.0}
    assert other[:, :24].min() >= 0.0 and other[:, :24].max() <= 1.0
    assert np.abs(other[:, 24]).max() <= 5.0 and np.isfinite(other).all()


@pytest.mark.parametrize("rows,cols,ranges", [
    (3, 40, [(2, 38), (0, 40), (1, 39)]),      # 120 cells, columns straddle a 32-bit word
    (2, 64, [(0, 64), (5, 60)]),               # 128 cells: the record's limit
    (1, 33, [(0, 33)]),                        # one row, just past the fast kernel's 32 columns
])
def test_wide_boards_take_the_generic_kernel_bit_exact(rows, cols, ranges):
    """Boards wider than 32 columns run features_kernel (16 lanes per state over the columns, fp64 arithmetic) instead of
    the bit-parallel features_fast_kernel. Same bit-exact bar, against the numpy oracle and its plain-C twin."""
    import ctypes as C
    from pathlib import Path

    from alphatriangle_b200.features import _run

    env = at.EnvConfig(ROWS=rows, COLS=cols, PLAYABLE_RANGE_PER_ROW=ranges, NUM_SHAPE_SLOTS=4)
    for channels, dim in ((1, 38), (2, 45), (1, 7)):
        mc = at.ModelConfig(GRID_INPUT_CHANNELS=channels, OTHER_NN_INPUT_FEATURES_DIM=dim)
        states = synth_states(333, env, seed=17 + cols) + edge_states(env)
        packed = pack_states(states, env)
        grid, other = _run(packed, env, channels, dim, torch.device("cuda:0"))
        g0, o0 = features_ref.extract_batch(states, mc)
        assert np.array_equal(bits(grid), bits(g0))
        assert np.array_equal(bits(other), bits(o0))
        lib = C.CDLL(str(Path(__file__).resolve().parents[1] / "oracle" / "liboracle_features.so"))
        gc, oc = np.empty_like(g0), np.empty_like(o0)
        assert lib.at_oracle_features(C.c_void_p(packed.ctypes.data), C.c_int64(len(states)), rows, cols, 4, channels, dim,
                                      C.c_void_p(gc.ctypes.data), C.c_void_p(oc.ctypes.data)) == 0
        assert np.array_equal(bits(grid), bits(gc)) and np.array_equal(bits(other), bits(oc))
