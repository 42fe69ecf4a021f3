"""Import shim for `trimcts` (needed because alphatriangle/config/__init__.py:3 imports
mcts_config.py, which imports SearchConfiguration). Fields: config/mcts_config.py:69-77."""

from pydantic import BaseModel


class SearchConfiguration(BaseModel):
    max_simulations: int = 64
    max_depth: int = 8
    cpuct: float = 1.0
    dirichlet_alpha: float = 0.3
    dirichlet_epsilon: float = 0.25
    discount: float = 1.0
    mcts_batch_size: int = 32


class AlphaZeroNetworkInterface:  # typing only
    pass
