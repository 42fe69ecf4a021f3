"""Import shim, NOT an engine: lets `import alphatriangle` (the unmodified reference under
/root/reference) succeed in a container without the compiled `trianglengin` package.
Only what the reference's nn / features / config modules import at module level is here
(features/extractor.py:7, nn/network.py:11, nn/model.py:9, config/__init__.py:1).
Field names follow the reference's tests/conftest.py:36-41; defaults are upstream's."""

from pydantic import BaseModel, Field


class EnvConfig(BaseModel):
    ROWS: int = Field(default=8, gt=0)
    COLS: int = Field(default=15, gt=0)
    PLAYABLE_RANGE_PER_ROW: list[tuple[int, int]] = Field(
        default=[(3, 12), (2, 13), (1, 14), (0, 15), (0, 15), (1, 14), (2, 13), (3, 12)]
    )
    NUM_SHAPE_SLOTS: int = Field(default=3, gt=0)


class GameState:  # typing only
    pass


class Shape:  # typing only
    pass
