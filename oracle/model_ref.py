"""PyTorch-CPU fp32 restatement of AlphaTriangleNet.forward and of the evaluate
post-processing (TEST INFRASTRUCTURE). Functional: it reads tensors out of a
reference-keyed state_dict instead of building modules.

  forward            <- nn/model.py:268-297 (module tree built at :116-266)
  _conv_unit         <- nn/model.py:15-38   (Conv2d -> BatchNorm2d(eval) -> activation)
  _res_block         <- nn/model.py:41-60
  _encoder_layer     <- torch.nn.TransformerEncoderLayer, norm_first=True, eval mode:
                        x + _sa_block(norm1(x)); x + _ff_block(norm2(x))  (math path, no mask)
  positional add     <- nn/model.py:63-106
  expected_value     <- nn/network.py:157-162, support at :61-67
  evaluate_batch     <- nn/network.py:242-318
"""

from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

from . import features_ref

_ACT = {
    "ReLU": F.relu,
    "GELU": F.gelu,
    "SiLU": F.silu,
    "Tanh": torch.tanh,
    "Sigmoid": torch.sigmoid,
}


def _pair(v):
    return (int(v[0]), int(v[1])) if isinstance(v, (tuple, list)) else (int(v), int(v))


def _bn(x, sd, prefix):
    return F.batch_norm(x, sd[prefix + ".running_mean"], sd[prefix + ".running_var"], sd[prefix + ".weight"],
                        sd[prefix + ".bias"], training=False, eps=1e-5)


def _conv_unit(x, sd, conv, bn, stride, pad, use_bn, act):
    x = F.conv2d(x, sd[conv + ".weight"], None if use_bn else sd[conv + ".bias"], stride=stride, padding=pad)
    if use_bn:
        x = _bn(x, sd, bn)
    return act(x)


def _res_block(x, sd, base, use_bn, act):
    out = _conv_unit(x, sd, base + ".conv1.0", base + ".conv1.1", 1, 1, use_bn, act)
    out = F.conv2d(out, sd[base + ".conv2.weight"], None if use_bn else sd[base + ".conv2.bias"], stride=1, padding=1)
    if use_bn:
        out = _bn(out, sd, base + ".bn2")
    return act(out + x)


def _encoder_layer(x, sd, base, heads, act):
    L, B, d = x.shape
    dh = d // heads
    h = F.layer_norm(x, (d,), sd[base + ".norm1.weight"], sd[base + ".norm1.bias"], 1e-5)
    qkv = F.linear(h, sd[base + ".self_attn.in_proj_weight"], sd[base + ".self_attn.in_proj_bias"])
    q, k, v = qkv.chunk(3, dim=-1)
    q = q.reshape(L, B * heads, dh).transpose(0, 1)
    k = k.reshape(L, B * heads, dh).transpose(0, 1)
    v = v.reshape(L, B * heads, dh).transpose(0, 1)
    att = torch.softmax(torch.bmm(q / math.sqrt(dh), k.transpose(1, 2)), dim=-1)
    o = torch.bmm(att, v).transpose(0, 1).reshape(L, B, d)
    x = x + F.linear(o, sd[base + ".self_attn.out_proj.weight"], sd[base + ".self_attn.out_proj.bias"])
    h = F.layer_norm(x, (d,), sd[base + ".norm2.weight"], sd[base + ".norm2.bias"], 1e-5)
    ff = F.linear(act(F.linear(h, sd[base + ".linear1.weight"], sd[base + ".linear1.bias"])),
                  sd[base + ".linear2.weight"], sd[base + ".linear2.bias"])
    return x + ff


def _mlp(x, sd, name, n_hidden, use_bn, act, has_out, taps=None):
    q = 0
    for _ in range(n_hidden):
        x = F.linear(x, sd[f"{name}.{q}.weight"], sd[f"{name}.{q}.bias"])
        if use_bn:
            x = _bn(x, sd, f"{name}.{q + 1}")
        x = act(x)
        if taps is not None:
            taps.append(x)
        q += 3 if use_bn else 2
    if has_out:
        x = F.linear(x, sd[f"{name}.{q}.weight"], sd[f"{name}.{q}.bias"])
    return x


@torch.no_grad()
def forward(sd: dict, mc, grid: torch.Tensor, other: torch.Tensor, taps: list | None = None):
    """(policy_logits [B,A], value_logits [B,atoms]) in fp32 on the CPU. `taps`, if given, collects the
    activation after each shared_fc layer."""
    act = _ACT[mc.ACTIVATION_FUNCTION]
    use_bn = mc.USE_BATCH_NORM
    x = grid
    for i in range(len(mc.CONV_FILTERS)):
        pad = mc.CONV_PADDING[i]
        pad = pad if isinstance(pad, str) else _pair(pad)
        x = _conv_unit(x, sd, f"conv_body.{i}.0", f"conv_body.{i}.1", _pair(mc.CONV_STRIDES[i]), pad, use_bn, act)
    idx = 0
    if mc.NUM_RESIDUAL_BLOCKS > 0:
        if x.shape[1] != mc.RESIDUAL_BLOCK_FILTERS:
            x = _conv_unit(x, sd, "res_body.0.0", "res_body.0.1", 1, 0, use_bn, act)
            idx = 1
        for _ in range(mc.NUM_RESIDUAL_BLOCKS):
            x = _res_block(x, sd, f"res_body.{idx}", use_bn, act)
            idx += 1
    if mc.USE_TRANSFORMER and mc.TRANSFORMER_LAYERS > 0:
        if "input_proj.weight" in sd:
            x = F.conv2d(x, sd["input_proj.weight"], sd["input_proj.bias"])
        seq = x.flatten(2).permute(2, 0, 1)
        seq = seq + sd["pos_encoder.pe"][: seq.size(0)]
        tact = F.relu if mc.ACTIVATION_FUNCTION == "ReLU" else F.gelu
        for l in range(mc.TRANSFORMER_LAYERS):
            seq = _encoder_layer(seq, sd, f"transformer_body.layers.{l}", mc.TRANSFORMER_HEADS, tact)
        d = seq.shape[-1]
        seq = F.layer_norm(seq, (d,), sd["transformer_body.norm.weight"], sd["transformer_body.norm.bias"], 1e-5)
        flat = seq.permute(1, 0, 2).flatten(1)
    else:
        flat = x.reshape(x.size(0), -1)
    trunk = _mlp(torch.cat([flat, other], dim=1), sd, "shared_fc", len(mc.FC_DIMS_SHARED), use_bn, act, False, taps)
    policy = _mlp(trunk, sd, "policy_head", len(mc.POLICY_HEAD_DIMS), use_bn, act, True)
    value = _mlp(trunk, sd, "value_head", len(mc.VALUE_HEAD_DIMS), use_bn, act, True)
    return policy, value


def expected_value(value_logits: torch.Tensor, mc) -> torch.Tensor:
    support = torch.linspace(mc.VALUE_MIN, mc.VALUE_MAX, mc.NUM_VALUE_ATOMS)
    probs = F.softmax(value_logits, dim=1)
    return torch.sum(probs * support.expand_as(probs), dim=1)


@torch.no_grad()
def evaluate_tensors(sd, mc, grid: np.ndarray, other: np.ndarray):
    """features -> (policy probs [B,A] float32, expected value [B] float32), the tensor-level result."""
    pl, vl = forward(sd, mc, torch.from_numpy(grid), torch.from_numpy(other))
    return F.softmax(pl, dim=1).numpy(), expected_value(vl, mc).numpy()


class NetworkEvaluationError(Exception):
    pass


def evaluate_batch(sd, mc, states) -> list:
    """The whole reference path on the CPU: per-state features, stack, forward, softmax, dicts."""
    if not states:
        return []
    grid, other = features_ref.extract_batch(states, mc)
    if not np.all(np.isfinite(grid)):
        raise NetworkEvaluationError("Non-finite values found in batched input grid_tensor")
    if not np.all(np.isfinite(other)):
        raise NetworkEvaluationError("Non-finite values found in batched input other_features_tensor")
    probs, values = evaluate_tensors(sd, mc, grid, other)
    results = []
    for i in range(len(states)):
        p = np.maximum(probs[i], 0)
        s = np.sum(p)
        if abs(s - 1.0) > 1e-5:
            if s <= 1e-9:
                valid = states[i].valid_actions()
                if not valid:
                    raise NetworkEvaluationError(f"Policy probability sum is near zero ({s}) for batch sample {i}")
                p = np.zeros_like(p)
                for a in valid:
                    if 0 <= a < len(p):
                        p[a] = 1.0 / len(valid)
            else:
                p = p / s
        results.append(({a: float(v) for a, v in enumerate(p)}, float(values[i])))
    return results
