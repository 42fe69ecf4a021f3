"""Drop-in for `alphatriangle.features.extract_state_features` (features/extractor.py:150-171).

The self-play worker stores the StateType of every move in the replay buffer
(alphatriangle/rl/self_play/worker.py:357). This module serves that caller from
the same CUDA kernel the evaluator uses: states are packed on the host, the
kernel writes grid / other_features, and the result comes back as the float32
numpy arrays the reference returns -- bit for bit. `extract_features_batch`
amortises the launch and the copies over many states.
"""

from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .packing import pack_states

StateType = dict  # {"grid": float32 (C,H,W), "other_features": float32 (D,)} -- alphatriangle/utils/types.py:8-15


def _run(packed: np.ndarray, env_config, grid_channels: int, other_dim: int, device: torch.device):
    lib = _lib.load()
    b = packed.shape[0]
    rows, cols = int(env_config.ROWS), int(env_config.COLS)
    grid = torch.empty((b, grid_channels, rows, cols), dtype=torch.float32, device=device)
    other = torch.empty((b, other_dim), dtype=torch.float32, device=device)
    if b == 0:
        return grid.cpu().numpy(), other.cpu().numpy()
    d_packed = torch.from_numpy(packed).to(device)
    env = _lib.env_desc(env_config)
    with torch.cuda.device(device):
        stream = C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        _lib.check(
            lib.at_extract_features(
                C.byref(env), grid_channels, other_dim, C.c_void_p(d_packed.data_ptr()), b, C.c_void_p(grid.data_ptr()),
                C.c_void_p(other.data_ptr()), stream,
            ),
            "at_extract_features",
        )
    return grid.cpu().numpy(), other.cpu().numpy()


def extract_features_batch(game_states, model_config, device: torch.device | str | None = None) -> list[StateType]:
    """StateType for each state, computed in one kernel launch."""
    if not game_states:
        return []
    device = torch.device(device if device is not None else "cuda")
    if device.type != "cuda":
        raise _lib.LibraryError("feature extraction runs on CUDA only; there is no CPU path")
    if device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    env_config = game_states[0].env_config
    packed = pack_states(game_states, env_config)
    grid, other = _run(packed, env_config, model_config.GRID_INPUT_CHANNELS, model_config.OTHER_NN_INPUT_FEATURES_DIM, device)
    return [{"grid": grid[i], "other_features": other[i]} for i in range(len(game_states))]


def extract_state_features(game_state, model_config, device: torch.device | str | None = None) -> StateType:
    """One state -> {"grid", "other_features"}; same values as the reference extractor."""
    return extract_features_batch([game_state], model_config, device)[0]
