"""Parameter container with the reference's `state_dict` key set.

`AlphaTriangleNet` here exists for two reasons only:
  * `NeuralNetwork.get_weights()/set_weights()` must produce and accept exactly
    the key set, shapes and dtypes of the reference model
    (alphatriangle/nn/model.py:116-266 builds them; network.py:320-336 moves them),
    including BatchNorm running statistics, `num_batches_tracked` and the
    non-trainable `pos_encoder.pe` buffer;
  * `Trainer` reaches through `nn.model` for its own forward/backward
    (alphatriangle/rl/core/trainer.py:34,230). `forward` below serves that
    caller with autograd; it is NOT on the evaluation path -- `evaluate_state`
    / `evaluate_batch` run the CUDA engine on a packed copy of these tensors
    and fail if the engine is missing.
"""

from __future__ import annotations

import math

import torch
import torch.nn as nn

from .config import ModelConfig


def _pair(v) -> tuple[int, int]:
    if isinstance(v, (tuple, list)):
        return int(v[0]), int(v[1])
    return int(v), int(v)


def _conv_unit(cin: int, cout: int, k, stride, pad, bn: bool, act: type[nn.Module]) -> nn.Sequential:
    mods: list[nn.Module] = [nn.Conv2d(cin, cout, k, stride, pad, bias=not bn)]
    if bn:
        mods.append(nn.BatchNorm2d(cout))
    mods.append(act())
    return nn.Sequential(*mods)


class ResidualBlock(nn.Module):
    """act(bn2(conv2(conv1(x))) + x); submodule names follow model.py:41-60."""

    def __init__(self, ch: int, bn: bool, act: type[nn.Module]):
        super().__init__()
        self.conv1 = _conv_unit(ch, ch, 3, 1, 1, bn, act)
        self.conv2 = nn.Conv2d(ch, ch, 3, 1, 1, bias=not bn)
        self.bn2 = nn.BatchNorm2d(ch) if bn else nn.Identity()
        self.activation = act()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.activation(self.bn2(self.conv2(self.conv1(x))) + x)


def sinusoid_table(max_len: int, d_model: int) -> torch.Tensor:
    """pe[l, 2i] = sin(l * w_i), pe[l, 2i+1] = cos(l * w_i), w_i = exp(-2i ln(1e4)/d) (model.py:72-81)."""
    pos = torch.arange(max_len).unsqueeze(1)
    freq = torch.exp(torch.arange(0, d_model, 2) * (-math.log(10000.0) / d_model))
    pe = torch.zeros(max_len, d_model)
    pe[:, 0::2] = torch.sin(pos * freq)
    if d_model > 1:
        pe[:, 1::2] = torch.cos(pos * freq)[:, : d_model // 2]
    return pe.unsqueeze(1)


class PositionalEncoding(nn.Module):
    def __init__(self, d_model: int, dropout: float = 0.1, max_len: int = 5000):
        super().__init__()
        if d_model <= 0:
            raise ValueError("d_model must be positive for PositionalEncoding")
        self.dropout = nn.Dropout(p=dropout)
        self.register_buffer("pe", sinusoid_table(max_len, d_model))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.shape[0] > self.pe.shape[0]:
            raise ValueError(f"Input sequence length {x.shape[0]} exceeds max_len {self.pe.shape[0]} of PositionalEncoding")
        return self.dropout(x + self.pe[: x.size(0)])


def _mlp(in_dim: int, hidden: list[int], bn: bool, act: type[nn.Module], out_dim: int | None) -> tuple[nn.Sequential, int]:
    mods: list[nn.Module] = []
    for h in hidden:
        mods.append(nn.Linear(in_dim, h))
        if bn:
            mods.append(nn.BatchNorm1d(h))
        mods.append(act())
        in_dim = h
    if out_dim is not None:
        mods.append(nn.Linear(in_dim, out_dim))
        in_dim = out_dim
    return nn.Sequential(*mods), in_dim


class AlphaTriangleNet(nn.Module):
    def __init__(self, model_config: ModelConfig, env_config):
        super().__init__()
        mc = model_config
        self.model_config = mc
        self.env_config = env_config
        rows, cols = int(env_config.ROWS), int(env_config.COLS)
        self.action_dim = int(env_config.NUM_SHAPE_SLOTS * rows * cols)
        act: type[nn.Module] = getattr(nn, mc.ACTIVATION_FUNCTION)
        bn = mc.USE_BATCH_NORM

        ch = mc.GRID_INPUT_CHANNELS
        h, w = rows, cols
        convs = []
        for i, cout in enumerate(mc.CONV_FILTERS):
            convs.append(_conv_unit(ch, cout, mc.CONV_KERNEL_SIZES[i], mc.CONV_STRIDES[i], mc.CONV_PADDING[i], bn, act))
            ch = cout
        self.conv_body = nn.Sequential(*convs)

        res: list[nn.Module] = []
        if mc.NUM_RESIDUAL_BLOCKS > 0:
            if ch != mc.RESIDUAL_BLOCK_FILTERS:
                res.append(_conv_unit(ch, mc.RESIDUAL_BLOCK_FILTERS, 1, 1, 0, bn, act))
                ch = mc.RESIDUAL_BLOCK_FILTERS
            res.extend(ResidualBlock(ch, bn, act) for _ in range(mc.NUM_RESIDUAL_BLOCKS))
        self.res_body = nn.Sequential(*res)
        self.cnn_output_channels = ch

        with torch.no_grad():
            probe = self.res_body(self.conv_body(torch.zeros(1, mc.GRID_INPUT_CHANNELS, rows, cols)))
        h, w = int(probe.shape[2]), int(probe.shape[3])
        self.out_hw = (h, w)

        self.input_proj: nn.Module = nn.Identity()
        self.pos_encoder = None
        self.transformer_body = None
        if mc.transformer_on:
            d = mc.TRANSFORMER_DIM
            if ch != d:
                self.input_proj = nn.Conv2d(ch, d, kernel_size=1)
            self.pos_encoder = PositionalEncoding(d, dropout=0.1)
            layer = nn.TransformerEncoderLayer(
                d_model=d,
                nhead=mc.TRANSFORMER_HEADS,
                dim_feedforward=mc.TRANSFORMER_FC_DIM,
                activation=mc.ACTIVATION_FUNCTION.lower(),
                batch_first=False,
                norm_first=True,
            )
            self.transformer_body = nn.TransformerEncoder(
                layer, num_layers=mc.TRANSFORMER_LAYERS, norm=nn.LayerNorm(d), enable_nested_tensor=False
            )
            flat = h * w * d
        else:
            flat = ch * h * w
        self.flat_features = flat

        self.shared_fc, trunk = _mlp(flat + mc.OTHER_NN_INPUT_FEATURES_DIM, list(mc.FC_DIMS_SHARED), bn, act, None)
        self.policy_head, _ = _mlp(trunk, list(mc.POLICY_HEAD_DIMS), bn, act, self.action_dim)
        self.value_head, _ = _mlp(trunk, list(mc.VALUE_HEAD_DIMS), bn, act, mc.NUM_VALUE_ATOMS)

    def forward(self, grid_state: torch.Tensor, other_features: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
        """Autograd forward for the Trainer (model.py:268-297). Not used by evaluate_*."""
        x = self.res_body(self.conv_body(grid_state))
        if self.transformer_body is not None and self.pos_encoder is not None:
            x = self.input_proj(x)
            seq = self.pos_encoder(x.flatten(2).permute(2, 0, 1))
            flat = self.transformer_body(seq).permute(1, 0, 2).flatten(1)
        else:
            flat = x.reshape(x.size(0), -1)
        trunk = self.shared_fc(torch.cat([flat, other_features], dim=1))
        return self.policy_head(trunk), self.value_head(trunk)
