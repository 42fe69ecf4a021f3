"""Drop-in for `alphatriangle.nn.NeuralNetwork` (alphatriangle/nn/network.py:32-336).

Same constructor, methods, public attributes and error convention, so trimcts,
the self-play worker and the trainer can hold this object instead of the
reference's. What differs is where the work happens: states are reduced to
80-byte records on the host, and feature extraction, the network forward and the
softmax / expected-value epilogues run as sm_100a kernels behind the C ABI. No
torch operator runs on the evaluation path, and nothing falls back to the CPU.
"""

from __future__ import annotations

import logging
from typing import Any

import numpy as np
import torch

from . import _lib
from .config import ModelConfig
from .engine import Engine
from .flat import FlatWeights
from .net import AlphaTriangleNet
from . import packing
from .packing import RECORD_BYTES, pack_states

logger = logging.getLogger(__name__)

PolicyValueOutput = tuple[dict[int, float], float]


class NetworkEvaluationError(Exception):
    """Raised for every failure of evaluate_state / evaluate_batch (network.py:26-29)."""


_PRECISIONS = {"fp32": _lib.PREC_FP32, "bf16": _lib.PREC_BF16, "f16": _lib.PREC_F16}


class NeuralNetwork:
    """Conforms to trimcts.AlphaZeroNetworkInterface, evaluating on one B200.

    Extra keyword arguments (not in the reference): `precision` -- "fp32" (parity
    path, max-abs 1e-4 against the reference), "f16" / "bf16" (the tcgen05 path with
    fp16 or bf16 operands, fp32 accumulation: policy KL <= 1e-3; fp16's 11-bit mantissa
    holds that bound also for near-deterministic policies, bf16's 8 bits hold it up to
    policies of about two nats) or "auto": fp16 when the
    tcgen05 path covers the configuration and the network has normalisation layers
    (activations then stay far inside fp16's range; conversions saturate at 65504),
    bf16 when it has none, fp32 when the tcgen05 path does not cover the configuration.
    `max_batch` -- capacity of the engine's staging buffers (larger batches are
    evaluated in slices).
    """

    def __init__(
        self,
        model_config: ModelConfig,
        env_config: Any,
        train_config: Any = None,
        device: torch.device | str | None = None,
        *,
        precision: str = "auto",
        max_batch: int = 4096,
    ):
        self.model_config = model_config
        self.env_config = env_config
        self.train_config = train_config
        device = torch.device(device if device is not None else "cuda")
        if device.type != "cuda":
            raise _lib.LibraryError(f"NeuralNetwork(device={device}): this implementation runs on CUDA sm_100a only")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = device
        self._orig_model = AlphaTriangleNet(model_config, env_config).to(device)
        self.model = self._orig_model  # no torch.compile on this path (COMPILE_MODEL is ignored)
        self.action_dim = int(env_config.NUM_SHAPE_SLOTS * env_config.ROWS * env_config.COLS)
        self.model.eval()

        self.num_atoms = model_config.NUM_VALUE_ATOMS
        self.v_min = model_config.VALUE_MIN
        self.v_max = model_config.VALUE_MAX
        self.delta_z = (self.v_max - self.v_min) / (self.num_atoms - 1)
        self.support = torch.linspace(self.v_min, self.v_max, self.num_atoms, device=self.device)

        self.engine = Engine(model_config, env_config, device, max_batch)
        if precision == "auto":
            if not self.engine.supports(_lib.PREC_BF16):
                self.precision = _lib.PREC_FP32
                # not silent: this configuration runs 20-30x slower than one the tcgen05 kernels cover
                logger.warning("precision='auto': this ModelConfig is outside the tcgen05 path (%s); using the fp32 CUDA-core path",
                               _lib.last_error())
            else:
                self.precision = _lib.PREC_F16 if model_config.USE_BATCH_NORM else _lib.PREC_BF16
        elif precision in _PRECISIONS:
            self.precision = _PRECISIONS[precision]
            if not self.engine.supports(self.precision):
                raise _lib.LibraryError(f"precision {precision!r} unavailable: {_lib.last_error()}")
        else:
            raise ValueError(f"precision must be 'auto', 'fp32', 'bf16' or 'f16', got {precision!r}")
        self._action_keys = tuple(range(self.action_dim))
        # parameters and float buffers become views into one flat buffer in the engine's image order (flat.py): in-place
        # updates (optimiser steps, load_state_dict, an NCCL receive) write the engine's input directly
        self.flat = FlatWeights(self._orig_model, self.engine.table, self.engine.image_floats, self.device)
        self._stamp: tuple | None = None
        self._weights_version = 0
        self._sync_engine()

    # ------------------------------------------------------------------ weights
    def _module_stamp(self) -> tuple:
        return self.flat.stamp()

    def _sync_engine(self, force: bool = False) -> None:
        """Re-fold the weights into the engine if any tensor was written since the last fold (trainer steps mutate them in
        place; the check is the tensors' version counters, ~5 us). Anything that REPLACES parameter storage instead of
        writing it (`model.to(...)`, `.half()`) must be followed by `refresh()`."""
        stamp = self.flat.stamp()
        if not force and stamp == self._stamp:
            return
        if not self.flat.aliased():
            self.flat.adopt()
        self._weights_version += 1
        self.engine.load_image(self.flat.image, self._weights_version)
        self._stamp = self.flat.stamp()

    def refresh(self) -> None:
        """Re-adopt the module's tensors into the flat buffer and re-fold, whatever happened to them."""
        self._sync_engine(force=True)

    def get_weights(self) -> dict[str, torch.Tensor]:
        """state_dict of the underlying module, on the CPU (network.py:320-323)."""
        return {k: v.cpu() for k, v in self._orig_model.state_dict().items()}

    def set_weights(self, weights: dict[str, torch.Tensor]) -> None:
        """load_state_dict into the underlying module, then refresh the engine's folded copy (network.py:325-336)."""
        try:
            on_device = {k: v.to(self.device) for k, v in weights.items()}
            self._orig_model.load_state_dict(on_device)
            self.model.eval()
            self._sync_engine(force=True)
            logger.debug("NN weights set successfully into underlying model.")
        except Exception as e:
            logger.error(f"Error setting weights on NN instance: {e}", exc_info=True)
            raise

    # ---------------------------------------------------------------- features
    def _pack(self, states) -> np.ndarray:
        return pack_states(states, self.env_config)

    def _features_on_device(self, states) -> tuple[torch.Tensor, torch.Tensor]:
        packed = torch.from_numpy(self._pack(states)).to(self.device)
        return self.engine.extract_features(
            packed, self.model_config.GRID_INPUT_CHANNELS, self.model_config.OTHER_NN_INPUT_FEATURES_DIM
        )

    def _state_to_tensors(self, state) -> tuple[torch.Tensor, torch.Tensor]:
        """Feature tensors of one state on the device (network.py:104-119)."""
        return self._features_on_device([state])

    def _batch_states_to_tensors(self, states) -> tuple[torch.Tensor, torch.Tensor]:
        """Batched feature tensors on the device (network.py:121-155); empty input gives empty tensors of the right rank."""
        if not states:
            mc, ec = self.model_config, self.env_config
            return (
                torch.empty((0, mc.GRID_INPUT_CHANNELS, ec.ROWS, ec.COLS), device=self.device),
                torch.empty((0, mc.OTHER_NN_INPUT_FEATURES_DIM), device=self.device),
            )
        return self._features_on_device(states)

    def _logits_to_expected_value(self, value_logits: torch.Tensor) -> torch.Tensor:
        """Kept for callers that hold logits (network.py:157-162); the evaluate path computes this in its epilogue kernel."""
        probs = torch.softmax(value_logits, dim=1)
        return torch.sum(probs * self.support.expand_as(probs), dim=1, keepdim=True)

    # -------------------------------------------------------------- evaluation
    @staticmethod
    def _raise_for_flags(flags: int, batched: bool, what: str) -> None:
        pre = "batched input " if batched else "input "
        if flags & _lib.FLAG_PIPELINE_TIMEOUT:
            raise NetworkEvaluationError(f"tcgen05 pipeline wait timed out (status {(flags >> 8) & 0xFF:#x}){what}")
        if flags & _lib.FLAG_NONFINITE_GRID:
            raise NetworkEvaluationError(f"Non-finite values found in {pre}grid_tensor{what}")
        if flags & _lib.FLAG_NONFINITE_OTHER:
            raise NetworkEvaluationError(f"Non-finite values found in {pre}other_features_tensor{what}")
        if flags & _lib.FLAG_NONFINITE_POLICY_LOGITS:
            raise NetworkEvaluationError(f"Non-finite policy_logits detected{what}.")
        if flags & _lib.FLAG_NONFINITE_VALUE_LOGITS:
            raise NetworkEvaluationError(f"Non-finite value_logits detected{what}.")
        if flags & _lib.FLAG_NONFINITE_POLICY_PROBS:
            raise NetworkEvaluationError(f"Non-finite policy probabilities AFTER softmax{what}.")

    def _finish_policy(self, probs_i: np.ndarray, state, label: str) -> np.ndarray:
        """Clamp / renormalise / uniform-over-valid fallback of network.py:281-305 for one row."""
        probs_i = np.maximum(probs_i, 0)
        prob_sum = np.sum(probs_i)
        if abs(prob_sum - 1.0) > 1e-5:
            logger.warning(f"Policy probabilities sum to {prob_sum:.6f} (not 1.0) for {label}. Re-normalizing.")
            if prob_sum <= 1e-9:
                valid_actions = state.valid_actions()
                if valid_actions:
                    probs_i = np.zeros_like(probs_i)
                    uniform = 1.0 / len(valid_actions)
                    for a in valid_actions:
                        if 0 <= a < len(probs_i):
                            probs_i[a] = uniform
                    logger.warning(f"Policy sum was zero for {label}, returning uniform over valid actions.")
                else:
                    raise NetworkEvaluationError(
                        f"Policy probability sum is near zero ({prob_sum}) for {label} with no valid actions. Cannot normalize."
                    )
            else:
                probs_i = probs_i / prob_sum
        return probs_i

    def _evaluate_rows(self, states) -> tuple[np.ndarray, np.ndarray, int]:
        self._sync_engine()
        packed = self._pack(states)
        cap = self.engine.max_batch
        if len(states) <= cap:
            return self.engine.evaluate_host(packed, self.precision)
        # more states than the engine's staging holds: chunks go through the two host slots, so one chunk's results
        # come down over PCIe while the next chunk's kernels run
        probs = np.empty((len(states), self.action_dim), dtype=np.float32)
        values = np.empty((len(states),), dtype=np.float32)
        flags = 0
        starts = list(range(0, len(states), cap))

        def collect(k: int) -> int:
            p, v, f = self.engine.wait_host(k & 1)
            lo = starts[k]
            probs[lo : lo + cap] = p
            values[lo : lo + cap] = v
            return f

        for k, lo in enumerate(starts):
            if k >= 2:
                flags |= collect(k - 2)
            self.engine.submit_host(packed[lo : lo + cap], self.precision, k & 1)
        for k in range(max(len(starts) - 2, 0), len(starts)):
            flags |= collect(k)
        return probs, values, flags

    def evaluate_state(self, state) -> PolicyValueOutput:
        """(policy dict over all actions, expected value) for one state (network.py:164-240)."""
        self.model.eval()
        try:
            probs, values, flags = self._evaluate_rows([state])
            self._raise_for_flags(flags, False, f" for state {state}")
            probs_i = self._finish_policy(probs[0], state, f"state {getattr(state, 'current_step', '?')}")
            return dict(zip(self._action_keys, probs_i.tolist())), float(values[0])
        except Exception as e:
            logger.error(f"Exception during single evaluation for state {state}: {e}", exc_info=True)
            raise NetworkEvaluationError(f"Evaluation failed for state {state}: {e}") from e

    def evaluate_batch(self, states) -> list[PolicyValueOutput]:
        """List of (policy dict, expected value), one per state (network.py:242-318). Empty input gives []."""
        if not states:
            return []
        self.model.eval()
        try:
            probs, values, flags = self._evaluate_rows(states)
            self._raise_for_flags(flags, True, " in batch evaluation")
            # dicts and tuples are built in C (csrc/at_pyhost.c: max(p, 0) per entry, cached int keys); rows whose
            # clipped sum is not within 5e-6 of 1 come back in `recheck` and take the reference's exact per-row rule
            results, recheck = packing._host().build_results(
                np.ascontiguousarray(probs, dtype=np.float32), np.ascontiguousarray(values, dtype=np.float32),
                len(states), self.action_dim, self._action_keys)
            for i in recheck:
                row = self._finish_policy(probs[i], states[i], f"sample {i}")
                results[i] = (dict(zip(self._action_keys, row.tolist())), results[i][1])
        except Exception as e:
            logger.error(f"Error during batch evaluation: {e}", exc_info=True)
            raise NetworkEvaluationError(f"Batch evaluation failed: {e}") from e
        return results

    # ---------------------------------------------------- tensor-level entry points
    def evaluate_packed(self, packed: torch.Tensor):
        """PackedState tensor uint8 [B,80] on the device -> (probs [B,A], value [B]) on the device, no host sync."""
        self._sync_engine()
        if packed.shape[0] > self.engine.max_batch and self.precision in _lib.TENSOR_CORE_PRECISIONS:
            raise ValueError(f"batch {packed.shape[0]} exceeds max_batch {self.engine.max_batch}")
        probs, value, _flags = self.engine.evaluate_packed(packed, self.precision)
        return probs, value

    def evaluate_features(self, grid: torch.Tensor, other: torch.Tensor):
        """Prepared feature tensors -> (probs, value) on the device; raises on non-finite inputs like the reference does."""
        self._sync_engine()
        try:
            probs, value, flags = self.engine.forward(grid, other, _lib.PREC_FP32)
            self._raise_for_flags(flags, True, " in batch evaluation")
        except NetworkEvaluationError:
            raise
        except Exception as e:
            raise NetworkEvaluationError(f"Batch evaluation failed: {e}") from e
        return probs, value


assert RECORD_BYTES == _lib.STATE_BYTES
