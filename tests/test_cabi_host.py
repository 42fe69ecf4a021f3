"""CPU-side checks of the C ABI: the library loads, exports every symbol the header
declares, and its weight-image table names exactly the reference state_dict's float keys.
No compute entry point is called here (no GPU in this tier)."""

import ctypes as C
import re
from pathlib import Path

import pytest
from golden_cases import CASES, case_configs

import alphatriangle_b200 as at
from alphatriangle_b200 import _lib

ROOT = Path(__file__).resolve().parents[1]


def test_library_exports_every_declared_symbol():
    header = (ROOT / "include" / "at_b200.h").read_text()
    declared = set(re.findall(r"\b(at_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    # the bring-up probes have their own header and library
    probe_header = (ROOT / "include" / "at_b200_probe.h").read_text()
    probe_declared = set(re.findall(r"\b(at_probe_[a-z0-9_]+)\s*\(", probe_header))
    assert probe_declared == set(_lib.PROBE_EXPORTS), probe_declared ^ set(_lib.PROBE_EXPORTS)
    probe = _lib.load_probe()
    assert all(hasattr(probe, name) for name in _lib.PROBE_EXPORTS)
    assert not any(name.startswith("at_probe") for name in _lib.EXPORTS)
    lib = C.CDLL(str(_lib.LIB_PATH))
    for name in declared:
        assert hasattr(lib, name), f"{name} not exported"
    assert _lib.load().at_abi_version() == 1


def test_struct_sizes_match_header():
    assert C.sizeof(_lib.EnvDesc) == 12
    # 3 + 1 + 1 + 8 + 16*3 + 2 + 5 + (1+4)*3 + 1 + 2(float) + 3 ints
    assert C.sizeof(_lib.ModelDesc) == 4 * (3 + 1 + 1 + 8 + 48 + 2 + 5 + 15 + 1 + 2 + 3)


@pytest.mark.parametrize("name", CASES)
def test_weight_table_names_the_reference_state_dict(name):
    mc_kwargs, env_kwargs, _ = case_configs(name)
    mc, env = at.ModelConfig(**mc_kwargs), at.EnvConfig(**env_kwargs)
    table = _lib.weight_table(mc, env)
    sd = at.AlphaTriangleNet(mc, env).state_dict()
    float_keys = {k for k, v in sd.items() if v.dtype.is_floating_point}
    assert {n for n, _, _ in table} == float_keys
    assert len(table) == len(float_keys)
    pos = 0
    for key, off, numel in table:
        assert off == pos and numel == sd[key].numel()
        pos += (numel + 3) & ~3  # entries start 16-byte aligned (include/at_b200.h)


def test_describe_rejects_bad_geometry():
    mc = at.ModelConfig()
    env = at.EnvConfig(ROWS=12, COLS=12, PLAYABLE_RANGE_PER_ROW=[(0, 12)] * 12, NUM_SHAPE_SLOTS=3)
    with pytest.raises(_lib.LibraryError, match="ROWS\\*COLS"):
        _lib.weight_table(mc, env)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setenv("AT_B200_LIB", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.LibraryError, match="no fallback"):
        _lib.load()
    monkeypatch.delenv("AT_B200_LIB")
    monkeypatch.setattr(_lib, "_lib", None)
    _lib.load()


def test_cpu_device_is_refused():
    import torch

    with pytest.raises(_lib.LibraryError, match="CUDA"):
        at.NeuralNetwork(at.ModelConfig(), at.EnvConfig(), None, torch.device("cpu"))
