"""Shape contract of the leaf-evaluation path.

`ModelConfig` keeps the reference's field names, defaults and validation errors
(reference: alphatriangle/config/model_config.py:7-157) because every tensor
dimension of the network is derived from it and callers construct it by keyword.
`EnvConfig` is `trianglengin.EnvConfig` when that package is importable; the
stand-in below carries the four fields the path reads (reference call sites:
alphatriangle/features/extractor.py:25,39,50 and alphatriangle/nn/model.py:123-125;
field names pinned by the reference's tests/conftest.py:36-41).
`TrainConfig` is accepted duck-typed by `NeuralNetwork`; only `COMPILE_MODEL`
(alphatriangle/config/train_config.py:89) is ever looked at, and it is ignored:
there is no tracing compiler on this path.
"""

from __future__ import annotations

from typing import Literal

from pydantic import BaseModel, Field, model_validator

try:  # the real engine, when the deployment has it
    from trianglengin import EnvConfig  # type: ignore  # noqa: F401

    HAVE_TRIANGLENGIN = True
except Exception:  # pragma: no cover - exercised in this container
    HAVE_TRIANGLENGIN = False

    class EnvConfig(BaseModel):  # type: ignore[no-redef]
        """Stand-in for trianglengin.EnvConfig (upstream defaults: 8x15 diamond, 3 slots)."""

        ROWS: int = Field(default=8, gt=0)
        COLS: int = Field(default=15, gt=0)
        PLAYABLE_RANGE_PER_ROW: list[tuple[int, int]] = Field(
            default=[(3, 12), (2, 13), (1, 14), (0, 15), (0, 15), (1, 14), (2, 13), (3, 12)]
        )
        NUM_SHAPE_SLOTS: int = Field(default=3, gt=0)

        @model_validator(mode="after")
        def _ranges(self) -> "EnvConfig":
            if len(self.PLAYABLE_RANGE_PER_ROW) != self.ROWS:
                raise ValueError("PLAYABLE_RANGE_PER_ROW length must equal ROWS.")
            for lo, hi in self.PLAYABLE_RANGE_PER_ROW:
                if not (0 <= lo <= hi <= self.COLS):
                    raise ValueError("PLAYABLE_RANGE_PER_ROW entries must satisfy 0 <= lo <= hi <= COLS.")
            return self


class ModelConfig(BaseModel):
    """Network dimensions; field-for-field the reference's ModelConfig."""

    GRID_INPUT_CHANNELS: int = Field(default=1, gt=0)

    CONV_FILTERS: list[int] = Field(default=[32, 64, 128])
    CONV_KERNEL_SIZES: list[int | tuple[int, int]] = Field(default=[3, 3, 3])
    CONV_STRIDES: list[int | tuple[int, int]] = Field(default=[1, 1, 1])
    CONV_PADDING: list[int | tuple[int, int] | str] = Field(default=[1, 1, 1])

    NUM_RESIDUAL_BLOCKS: int = Field(default=2, ge=0)
    RESIDUAL_BLOCK_FILTERS: int = Field(default=128, gt=0)

    USE_TRANSFORMER: bool = Field(default=True)
    TRANSFORMER_DIM: int = Field(default=128, gt=0)
    TRANSFORMER_HEADS: int = Field(default=4, gt=0)
    TRANSFORMER_LAYERS: int = Field(default=2, ge=0)
    TRANSFORMER_FC_DIM: int = Field(default=256, gt=0)

    FC_DIMS_SHARED: list[int] = Field(default=[128])
    POLICY_HEAD_DIMS: list[int] = Field(default=[128])

    NUM_VALUE_ATOMS: int = Field(default=51, gt=1)
    VALUE_MIN: float = Field(default=-10.0)
    VALUE_MAX: float = Field(default=10.0)
    VALUE_HEAD_DIMS: list[int] = Field(default=[128])

    ACTIVATION_FUNCTION: Literal["ReLU", "GELU", "SiLU", "Tanh", "Sigmoid"] = Field(default="ReLU")
    USE_BATCH_NORM: bool = Field(default=True)

    OTHER_NN_INPUT_FEATURES_DIM: int = Field(default=30, gt=0)

    @model_validator(mode="after")
    def _check(self) -> "ModelConfig":
        n = len(self.CONV_FILTERS)
        if not (len(self.CONV_KERNEL_SIZES) == n and len(self.CONV_STRIDES) == n and len(self.CONV_PADDING) == n):
            raise ValueError(
                "Lengths of CONV_FILTERS, CONV_KERNEL_SIZES, CONV_STRIDES, and CONV_PADDING must match."
            )
        if self.USE_TRANSFORMER and self.TRANSFORMER_LAYERS > 0:
            if self.TRANSFORMER_DIM % self.TRANSFORMER_HEADS != 0:
                raise ValueError(
                    f"TRANSFORMER_DIM ({self.TRANSFORMER_DIM}) must be divisible by "
                    f"TRANSFORMER_HEADS ({self.TRANSFORMER_HEADS})."
                )
        if self.VALUE_MIN >= self.VALUE_MAX:
            raise ValueError("VALUE_MIN must be strictly less than VALUE_MAX.")
        return self

    @property
    def transformer_on(self) -> bool:
        return bool(self.USE_TRANSFORMER and self.TRANSFORMER_LAYERS > 0)


class TrainConfig(BaseModel):
    """The two TrainConfig fields a `NeuralNetwork` constructor can meet on this path."""

    COMPILE_MODEL: bool = Field(default=False)
    DEVICE: str = Field(default="cuda")


def playable_mask(env_config) -> "list[list[bool]]":
    """death[r][c] is the complement of this: cells outside PLAYABLE_RANGE_PER_ROW are dead."""
    rows, cols = int(env_config.ROWS), int(env_config.COLS)
    out = [[False] * cols for _ in range(rows)]
    for r, (lo, hi) in enumerate(env_config.PLAYABLE_RANGE_PER_ROW):
        for c in range(int(lo), int(hi)):
            out[r][c] = True
    return out
