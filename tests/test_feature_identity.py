"""The feature kernel's fast path replaces the reference's float64 divisions of small integers by IEEE float32
divisions (csrc/at_features.cu). That is bit-exact iff float32(float64(a)/float64(b)) == float32(a)/float32(b); checked
here exhaustively over every operand a PackedState can produce, for the single, two-step and width quotients."""

import numpy as np


def _bits(x):
    return np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)


def test_single_division_identity():
    a = np.arange(-800, 1201, dtype=np.int64)[:, None]  # sums of int8 bbox coordinates, n_tris, heights, holes, step < 1000 ...
    b = np.arange(1, 2049, dtype=np.int64)[None, :]     # 5, 1000, rows, cols, 2*rows, 4*cols, playable cells, n_tris ...
    ref = (a.astype(np.float64) / b.astype(np.float64)).astype(np.float32)
    got = a.astype(np.float32) / b.astype(np.float32)
    assert np.array_equal(_bits(ref), _bits(got))


def test_two_step_division_identity():
    a = np.arange(-600, 1025, dtype=np.int64)[:, None]
    b = np.arange(1, 129, dtype=np.int64)[None, :]
    for c in range(1, 129):  # (a / b) / c as the reference computes mean(heights)/rows, (bump/(cols-1))/rows, ((min+max)/2)/rows
        ref = ((a.astype(np.float64) / b.astype(np.float64)) / np.float64(c)).astype(np.float32)
        got = a.astype(np.float32) / (b * c).astype(np.float32)
        assert np.array_equal(_bits(ref), _bits(got)), c


def test_effective_width_identity():
    w = np.arange(-254, 257, dtype=np.int64)[:, None]  # max_c - min_c + 1 for int8 coordinates
    cols = np.arange(1, 129, dtype=np.int64)[None, :]
    ref = ((w.astype(np.float64) * 0.75 + 0.25) / cols.astype(np.float64)).astype(np.float32)
    got = (3 * w + 1).astype(np.float32) / (4 * cols).astype(np.float32)
    assert np.array_equal(_bits(ref), _bits(got))
    clipped_ref = np.clip((w.astype(np.float64) * 0.75 + 0.25) / cols.astype(np.float64), 0, 1).astype(np.float32)
    assert np.array_equal(_bits(clipped_ref), _bits(np.clip(got, np.float32(0), np.float32(1))))
