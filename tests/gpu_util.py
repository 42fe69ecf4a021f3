"""Helpers shared by the -m gpu tests."""

import numpy as np
import torch
from golden_cases import case_configs, case_weights

import alphatriangle_b200 as at

_NETS: dict = {}


def load_case(name, golden_dir):
    mc_kwargs, env_kwargs, seed = case_configs(name)
    return at.ModelConfig(**mc_kwargs), at.EnvConfig(**env_kwargs), seed, np.load(golden_dir / f"{name}.npz")


def network(name, precision="fp32", max_batch=4096):
    """One NeuralNetwork per (case, precision), carrying the seeded weights the golden files were made with."""
    key = (name, precision, max_batch)
    if key not in _NETS:
        mc_kwargs, env_kwargs, seed = case_configs(name)
        mc, env = at.ModelConfig(**mc_kwargs), at.EnvConfig(**env_kwargs)
        nn = at.NeuralNetwork(mc, env, at.TrainConfig(), torch.device("cuda:0"), precision=precision, max_batch=max_batch)
        nn.set_weights(case_weights(name, nn.get_weights()))
        _NETS[key] = nn
    return _NETS[key]


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def kl_rows(p_ref, q):
    p = np.maximum(p_ref.astype(np.float64), 1e-30)
    q = np.maximum(q.astype(np.float64), 1e-30)
    return np.sum(p * (np.log(p) - np.log(q)), axis=1)
