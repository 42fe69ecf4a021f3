"""PackedState: the fixed 80-byte device record one game state is reduced to.

The record holds exactly what the reference's extractor reads through the
GameState accessor surface (alphatriangle/features/extractor.py:28-31,58-66,109,114)
and nothing derived from it: every division, clip and rounding of the feature
vector happens on the GPU (csrc/at_features.cu).

    offset  size  field
    0       16    occupied bitmask, bit r*COLS+c of a little-endian 128-bit word
    16      16    death bitmask, same indexing (read per state, as the reference does)
    32      4*8   shape slots: u8 present (truthiness of the Shape object),
                  u8 n_tris = len(shape.triangles), u8 n_up = #is_up, u8 0,
                  i8 min_r, i8 min_c, i8 max_r, i8 max_c  (= shape.bbox())
    64      8     game_score() as float64
    72      4     current_step as int32 (saturated; the feature clips at 1000 anyway)
    76      4     zero

Limits (rejected here with ValueError, never silently wrapped): ROWS*COLS <= 128,
NUM_SHAPE_SLOTS <= 4, n_tris <= 255, bbox coordinates within int8.
"""

from __future__ import annotations

import numpy as np

RECORD_BYTES = 80
MAX_CELLS = 128
MAX_SLOTS = 4

PACKED_DTYPE = np.dtype(
    {
        "names": ["occ", "death", "slots", "score", "step", "pad"],
        "formats": [("u1", 16), ("u1", 16), ("i1", (MAX_SLOTS, 8)), "<f8", "<i4", "<u4"],
        "offsets": [0, 16, 32, 64, 72, 76],
        "itemsize": RECORD_BYTES,
    }
)


def check_env(env_config) -> None:
    rows, cols, slots = int(env_config.ROWS), int(env_config.COLS), int(env_config.NUM_SHAPE_SLOTS)
    if rows <= 0 or cols <= 0 or rows * cols > MAX_CELLS:
        raise ValueError(f"PackedState supports 0 < ROWS*COLS <= {MAX_CELLS}, got {rows}x{cols}")
    if slots <= 0 or slots > MAX_SLOTS:
        raise ValueError(f"PackedState supports 1..{MAX_SLOTS} shape slots, got {slots}")


def _bits(mask: np.ndarray) -> np.ndarray:
    flat = np.ascontiguousarray(mask, dtype=bool).reshape(-1)
    out = np.zeros(16, dtype=np.uint8)
    packed = np.packbits(flat, bitorder="little")
    out[: packed.shape[0]] = packed
    return out


def pack_state_into(rec: np.ndarray, state, rows: int, cols: int, slots: int) -> None:
    """Fill one PACKED_DTYPE record from a GameState-like object."""
    grid = state.get_grid_data_np()
    occ = grid["occupied"]
    death = grid["death"]
    if occ.shape != (rows, cols) or death.shape != (rows, cols):
        raise ValueError(f"grid arrays have shape {occ.shape}/{death.shape}, expected {(rows, cols)}")
    rec["occ"] = _bits(occ)
    rec["death"] = _bits(death)
    shapes = state.get_shapes()
    if len(shapes) > slots:
        raise ValueError(f"state has {len(shapes)} shape slots, EnvConfig says {slots}")
    srec = rec["slots"]
    srec[...] = 0
    for i, shape in enumerate(shapes):
        if not shape:
            continue
        srec[i, 0] = 1
        tris = shape.triangles
        n = len(tris)
        if n == 0:
            continue
        if n > 255:
            raise ValueError(f"shape in slot {i} has {n} triangles; PackedState holds at most 255")
        ups = sum(1 for _, _, is_up in tris if is_up)
        min_r, min_c, max_r, max_c = shape.bbox()
        for v in (min_r, min_c, max_r, max_c):
            if not -128 <= int(v) <= 127:
                raise ValueError(f"shape bbox coordinate {v} outside int8")
        srec[i, 1] = np.uint8(n).view(np.int8)
        srec[i, 2] = np.uint8(ups).view(np.int8)
        srec[i, 4] = min_r
        srec[i, 5] = min_c
        srec[i, 6] = max_r
        srec[i, 7] = max_c
    rec["score"] = float(state.game_score())
    step = int(state.current_step)
    rec["step"] = max(-(2**31), min(2**31 - 1, step))
    rec["pad"] = 0


def _host():
    """The C host loops (csrc/at_pyhost.c), built next to libat_b200.so."""
    global _pyhost
    if _pyhost is None:
        try:
            from . import _pyhost as mod
        except ImportError as e:  # same rule as the CUDA library: no silent slow path
            raise ImportError(
                "alphatriangle_b200/_pyhost.so not found: build it with `python __graft_entry__.py build` (gcc, Python.h)"
            ) from e
        _pyhost = mod
    return _pyhost


_pyhost = None


def pack_states_py(states, env_config, out: np.ndarray | None = None) -> np.ndarray:
    """Pure-Python twin of the C packer: the readable statement of the record format, the packer for array types the C
    loop does not read directly (anything that is not a one-byte-per-cell array), and the one that words shape errors."""
    check_env(env_config)
    rows, cols, slots = int(env_config.ROWS), int(env_config.COLS), int(env_config.NUM_SHAPE_SLOTS)
    n = len(states)
    recs = _records(n, out)
    for i, state in enumerate(states):
        pack_state_into(recs[i], state, rows, cols, slots)
    return recs.view(np.uint8).reshape(n, RECORD_BYTES)


def _records(n: int, out: np.ndarray | None) -> np.ndarray:
    if out is None:
        return np.zeros(n, dtype=PACKED_DTYPE)
    if out.dtype != np.uint8 or out.shape != (n, RECORD_BYTES) or not out.flags.c_contiguous:
        raise ValueError("out must be a C-contiguous uint8 array of shape [len(states), 80]")
    return out.reshape(-1).view(PACKED_DTYPE)


def pack_states(states, env_config, out: np.ndarray | None = None) -> np.ndarray:
    """list[GameState] -> uint8 [B, 80] (a view of a PACKED_DTYPE array). The loop runs in C (csrc/at_pyhost.c); a batch
    holding a state whose grid arrays are not plain byte arrays of shape [ROWS, COLS] goes through pack_states_py."""
    check_env(env_config)
    rows, cols, slots = int(env_config.ROWS), int(env_config.COLS), int(env_config.NUM_SHAPE_SLOTS)
    n = len(states)
    recs = _records(n, out)
    done = _host().pack_states(states, rows, cols, slots, recs.view(np.uint8).reshape(-1))
    if done < 0:
        return pack_states_py(states, env_config, recs.view(np.uint8).reshape(n, RECORD_BYTES))
    return recs.view(np.uint8).reshape(n, RECORD_BYTES)


def unpack_masks(packed: np.ndarray, rows: int, cols: int) -> tuple[np.ndarray, np.ndarray]:
    """uint8 [B,80] -> (occupied, death) bool [B,rows,cols]; used by tests and the oracle."""
    recs = np.ascontiguousarray(packed).reshape(-1).view(PACKED_DTYPE)
    occ = np.unpackbits(recs["occ"], axis=1, bitorder="little")[:, : rows * cols]
    death = np.unpackbits(recs["death"], axis=1, bitorder="little")[:, : rows * cols]
    return occ.reshape(-1, rows, cols).astype(bool), death.reshape(-1, rows, cols).astype(bool)
