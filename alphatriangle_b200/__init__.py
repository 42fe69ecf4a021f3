"""B200-native MCTS leaf evaluation for AlphaTriangle.

Public surface mirrors the reference packages `alphatriangle.nn` and
`alphatriangle.features` for this one path; see DESIGN.md and INTEGRATION.md.
"""

from .batcher import LeafBatcher
from .config import EnvConfig, ModelConfig, TrainConfig
from .features import extract_features_batch, extract_state_features
from .net import AlphaTriangleNet
from .network import NetworkEvaluationError, NeuralNetwork
from .packing import pack_states

__all__ = [
    "AlphaTriangleNet",
    "EnvConfig",
    "LeafBatcher",
    "ModelConfig",
    "NetworkEvaluationError",
    "NeuralNetwork",
    "TrainConfig",
    "extract_features_batch",
    "extract_state_features",
    "pack_states",
]
