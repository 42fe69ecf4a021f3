"""numpy restatement of alphatriangle/features (TEST INFRASTRUCTURE).

Follows, function by function:
  column_heights   <- features/grid_features.py:7-19   (numba get_column_heights)
  count_holes      <- features/grid_features.py:22-33
  bumpiness        <- features/grid_features.py:36-42
  grid_state       <- features/extractor.py:33-46
  shape_features   <- features/extractor.py:48-85
  shape_availability <- features/extractor.py:87-90
  explicit_features  <- features/extractor.py:92-118
  combined_other   <- features/extractor.py:120-147
  extract_state_features <- features/extractor.py:150-171
Arithmetic is done the way the reference does it: Python floats / numpy float64
scalars, one rounding when a value is stored into a float32 array.
"""

from __future__ import annotations

import numpy as np


def column_heights(occupied: np.ndarray, death: np.ndarray, rows: int, cols: int) -> np.ndarray:
    heights = np.zeros(cols, dtype=np.int32)
    for c in range(cols):
        max_r = -1
        for r in range(rows):
            if occupied[r, c] and not death[r, c]:
                max_r = r
        heights[c] = max_r + 1
    return heights


def count_holes(occupied: np.ndarray, death: np.ndarray, heights: np.ndarray, cols: int) -> int:
    holes = 0
    for c in range(cols):
        for r in range(int(heights[c])):
            if not occupied[r, c] and not death[r, c]:
                holes += 1
    return holes


def bumpiness(heights: np.ndarray) -> float:
    total = 0.0
    for i in range(len(heights) - 1):
        total += abs(int(heights[i]) - int(heights[i + 1]))
    return total


def grid_state(occupied, death, rows, cols, channels) -> np.ndarray:
    g = np.zeros((channels, rows, cols), dtype=np.float32)
    g[0, occupied & ~death] = 1.0
    g[0, death] = -1.0
    return g


def shape_features(shapes, rows, cols, slots) -> np.ndarray:
    m = np.zeros((slots, 7), dtype=np.float32)
    for i, shape in enumerate(shapes):
        if shape and shape.triangles:
            n = len(shape.triangles)
            ups = sum(1 for _, _, is_up in shape.triangles if is_up)
            downs = n - ups
            min_r, min_c, max_r, max_c = shape.bbox()
            height = max_r - min_r + 1
            width_eff = (max_c - min_c + 1) * 0.75 + 0.25
            m[i, 0] = np.clip(n / 5.0, 0, 1)
            m[i, 1] = ups / n
            m[i, 2] = downs / n
            m[i, 3] = np.clip(height / rows, 0, 1)
            m[i, 4] = np.clip(width_eff / cols, 0, 1)
            m[i, 5] = np.clip(((min_r + max_r) / 2.0) / rows, 0, 1)
            m[i, 6] = np.clip(((min_c + max_c) / 2.0) / cols, 0, 1)
    return m.flatten()


def shape_availability(shapes) -> np.ndarray:
    return np.array([1.0 if s else 0.0 for s in shapes], dtype=np.float32)


def explicit_features(occupied, death, rows, cols, score, step) -> np.ndarray:
    f = np.zeros(6, dtype=np.float32)
    heights = column_heights(occupied, death, rows, cols)
    holes = count_holes(occupied, death, heights, cols)
    bump = bumpiness(heights)
    playable = np.sum(~death)
    with np.errstate(all="ignore"):
        f[0] = np.clip(score / 100.0, -5.0, 5.0)
        f[1] = np.mean(heights) / rows if rows > 0 else 0
        f[2] = np.max(heights) / rows if rows > 0 else 0
        f[3] = holes / playable if playable > 0 else 0
        f[4] = (bump / (cols - 1)) / rows if cols > 1 and rows > 0 else 0
        f[5] = np.clip(step / 1000.0, 0, 1)
    return np.nan_to_num(f, nan=0.0, posinf=0.0, neginf=0.0)


def combined_other(state, rows, cols, slots, other_dim) -> np.ndarray:
    g = state.get_grid_data_np()
    shapes = state.get_shapes()
    parts = [
        shape_features(shapes, rows, cols, slots),
        shape_availability(state.get_shapes()),
        explicit_features(g["occupied"], g["death"], rows, cols, state.game_score(), state.current_step),
    ]
    combined = np.concatenate(parts)
    if combined.shape[0] < other_dim:
        combined = np.concatenate([combined, np.zeros(other_dim - combined.shape[0], dtype=combined.dtype)])
    elif combined.shape[0] > other_dim:
        combined = combined[:other_dim]
    if not np.all(np.isfinite(combined)):
        combined = np.nan_to_num(combined, nan=0.0, posinf=0.0, neginf=0.0)
    return combined.astype(np.float32)


def extract_state_features(state, model_config) -> dict:
    env = state.env_config
    rows, cols, slots = int(env.ROWS), int(env.COLS), int(env.NUM_SHAPE_SLOTS)
    g = state.get_grid_data_np()
    return {
        "grid": grid_state(g["occupied"], g["death"], rows, cols, model_config.GRID_INPUT_CHANNELS),
        "other_features": combined_other(state, rows, cols, slots, model_config.OTHER_NN_INPUT_FEATURES_DIM),
    }


def extract_batch(states, model_config) -> tuple[np.ndarray, np.ndarray]:
    feats = [extract_state_features(s, model_config) for s in states]
    return np.stack([f["grid"] for f in feats]), np.stack([f["other_features"] for f in feats])
