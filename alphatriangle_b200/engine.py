"""Python handle on the native engine (at_engine_* in include/at_b200.h).

torch is used here for what the task allows it for: device memory (output
tensors, pinned staging) and streams. All arithmetic happens inside the library.
"""

from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .packing import RECORD_BYTES


def _stream_ptr(device: torch.device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptr(t: torch.Tensor) -> C.c_void_p:
    return C.c_void_p(t.data_ptr())


class Engine:
    def __init__(self, model_config, env_config, device: torch.device, max_batch: int):
        if device.type != "cuda":
            raise _lib.LibraryError(
                f"alphatriangle_b200 evaluates on a B200 only (device={device}); there is no CPU path."
            )
        self.lib = _lib.load()
        self.device = torch.device("cuda", device.index if device.index is not None else torch.cuda.current_device())
        self.max_batch = int(max_batch)
        self.desc = _lib.model_desc(model_config, env_config)
        handle = C.c_void_p()
        _lib.check(
            self.lib.at_engine_create(C.byref(self.desc), self.device.index, self.max_batch, C.byref(handle)),
            "at_engine_create",
        )
        self.handle = handle
        self.action_dim = int(self.lib.at_engine_action_dim(self.handle))
        self.image_floats = int(self.lib.at_engine_weight_floats(self.handle))
        self.table = self._read_table()
        # pinned staging for the host entry point
        self._h_states = torch.empty((self.max_batch, RECORD_BYTES), dtype=torch.uint8, pin_memory=True)
        self._h_policy = torch.empty((self.max_batch, self.action_dim), dtype=torch.float32, pin_memory=True)
        self._h_value = torch.empty((self.max_batch,), dtype=torch.float32, pin_memory=True)
        self._h_flags = torch.zeros((4,), dtype=torch.int32, pin_memory=True)
        # evaluate_host is what a search calls with a few dozen leaves: its numpy views and ctypes pointers are made once
        self._h_np = (self._h_states.numpy(), self._h_policy.numpy(), self._h_value.numpy(), self._h_flags.numpy())
        self._h_ptrs = (_ptr(self._h_states), _ptr(self._h_policy), _ptr(self._h_value), _ptr(self._h_flags))
        self._slots: list = [None, None]  # pinned staging of the pipelined host calls, allocated on first use
        self._slot_batch = [0, 0]
        self._d_flags = torch.zeros((4,), dtype=torch.int32, device=self.device)

    def _read_table(self) -> list[tuple[str, int, int]]:
        out = []
        name, off, numel = C.c_char_p(), C.c_int64(), C.c_int64()
        for i in range(int(self.lib.at_engine_weight_entries(self.handle))):
            _lib.check(
                self.lib.at_engine_weight_entry(self.handle, i, C.byref(name), C.byref(off), C.byref(numel)),
                "at_engine_weight_entry",
            )
            out.append((name.value.decode(), int(off.value), int(numel.value)))
        return out

    def close(self) -> None:
        if getattr(self, "handle", None):
            self.lib.at_engine_destroy(self.handle)
            self.handle = None

    def __del__(self):  # pragma: no cover - interpreter teardown order
        try:
            self.close()
        except Exception:
            pass

    # ---- weights -------------------------------------------------------------------
    def build_image(self, state_dict: dict[str, torch.Tensor]) -> torch.Tensor:
        """A weight image (device, fp32) from a reference-keyed state_dict: each float tensor at its table offset. For callers
        that hold loose tensors; NeuralNetwork keeps its parameters inside such an image already (flat.py) and never packs."""
        image = torch.zeros(self.image_floats, dtype=torch.float32, device=self.device)
        for name, off, numel in self.table:
            if name not in state_dict:
                raise KeyError(f"state_dict is missing {name!r}")
            t = state_dict[name]
            if t.numel() != numel:
                raise ValueError(f"{name}: expected {numel} elements, got {tuple(t.shape)}")
            image[off : off + numel].copy_(t.detach().reshape(-1))
        return image

    def _check_image(self, image: torch.Tensor) -> None:
        if image.device != self.device or image.dtype != torch.float32 or not image.is_contiguous():
            raise ValueError("weight image must be a contiguous float32 tensor on the engine's device")

    def load_image(self, image: torch.Tensor, version: int = 0) -> None:
        """Fold / repack `image` on the current stream; the next evaluate call reads the new set (it waits for the fold)."""
        self._check_image(image)
        _lib.check(
            self.lib.at_engine_load_weights(self.handle, _ptr(image), image.numel(), int(version), _stream_ptr(self.device)),
            "at_engine_load_weights",
        )
        self._last_image = image  # keep alive until the fold kernels queued on the stream have read it

    def publish_image(self, image: torch.Tensor, version: int = 0) -> None:
        """Fold / repack `image` on the current stream (a side stream) while evaluations on other streams keep reading the
        old set; they switch over at the first evaluate call after the fold has finished (at_engine_publish_weights)."""
        self._check_image(image)
        _lib.check(
            self.lib.at_engine_publish_weights(self.handle, _ptr(image), image.numel(), int(version), _stream_ptr(self.device)),
            "at_engine_publish_weights",
        )
        self._last_image = image

    def weight_version(self) -> int:
        return int(self.lib.at_engine_weight_version(self.handle))

    def supports(self, precision: int) -> bool:
        return bool(self.lib.at_engine_supports(self.handle, precision))

    def last_launches(self) -> int:
        return int(self.lib.at_engine_last_launches(self.handle))

    # ---- evaluation ------------------------------------------------------------------
    def forward(self, grid: torch.Tensor, other: torch.Tensor, precision: int = _lib.PREC_FP32):
        """Prepared features (device, fp32) -> (probs [B,A], value [B], flags int)."""
        b = grid.shape[0]
        grid = grid.to(device=self.device, dtype=torch.float32).contiguous()
        other = other.to(device=self.device, dtype=torch.float32).contiguous()
        probs = torch.empty((b, self.action_dim), dtype=torch.float32, device=self.device)
        value = torch.empty((b,), dtype=torch.float32, device=self.device)
        self._d_flags.zero_()
        _lib.check(
            self.lib.at_engine_forward(
                self.handle, precision, _ptr(grid), _ptr(other), b, _ptr(probs), _ptr(value), _ptr(self._d_flags),
                _stream_ptr(self.device),
            ),
            "at_engine_forward",
        )
        return probs, value, int(self._d_flags[0].item())

    def evaluate_packed(self, packed: torch.Tensor, precision: int, probs: torch.Tensor | None = None,
                        value: torch.Tensor | None = None, flags: torch.Tensor | None = None):
        """PackedState[B] on the device -> probs / value on the device. Asynchronous (no host sync)."""
        if packed.device != self.device or packed.dtype != torch.uint8 or packed.dim() != 2 or packed.shape[1] != RECORD_BYTES:
            raise ValueError("packed must be a uint8 [B,80] tensor on the engine's device")
        if not packed.is_contiguous():
            packed = packed.contiguous()
        b = packed.shape[0]
        if probs is None:
            probs = torch.empty((b, self.action_dim), dtype=torch.float32, device=self.device)
        if value is None:
            value = torch.empty((b,), dtype=torch.float32, device=self.device)
        if flags is None:
            flags = self._d_flags
            flags.zero_()
        _lib.check(
            self.lib.at_engine_evaluate_packed(
                self.handle, precision, _ptr(packed), b, _ptr(probs), _ptr(value), _ptr(flags), _stream_ptr(self.device)
            ),
            "at_engine_evaluate_packed",
        )
        return probs, value, flags

    def evaluate_host(self, packed: np.ndarray | torch.Tensor, precision: int):
        """Host records -> (probs ndarray [B,A], value ndarray [B], flags int); views of pinned staging, valid until the next call."""
        b = int(packed.shape[0])
        if b > self.max_batch:
            raise ValueError(f"batch {b} exceeds max_batch {self.max_batch}")
        states_np, policy_np, value_np, flags_np = self._h_np
        if isinstance(packed, torch.Tensor):
            self._h_states[:b].copy_(packed)
        else:
            states_np[:b] = packed
        p_states, p_policy, p_value, p_flags = self._h_ptrs
        rc = self.lib.at_engine_evaluate_host(self.handle, precision, p_states, b, p_policy, p_value, p_flags, _stream_ptr(self.device))
        if rc != _lib.AT_OK:
            _lib.check(rc, "at_engine_evaluate_host")
        return policy_np[:b], value_np[:b], int(flags_np[0])

    # ---- two batches in flight (at_engine_submit_host / at_engine_wait_host) ----
    def _slot_buffers(self, slot: int):
        if slot not in (0, 1):
            raise ValueError("slot must be 0 or 1")
        if self._slots[slot] is None:
            self._slots[slot] = (
                torch.empty((self.max_batch, RECORD_BYTES), dtype=torch.uint8, pin_memory=True),
                torch.empty((self.max_batch, self.action_dim), dtype=torch.float32, pin_memory=True),
                torch.empty((self.max_batch,), dtype=torch.float32, pin_memory=True),
                torch.zeros((4,), dtype=torch.int32, pin_memory=True),
            )
        return self._slots[slot]

    def submit_host(self, packed: np.ndarray | torch.Tensor, precision: int, slot: int) -> None:
        """Queue one batch of host records on `slot` (0 / 1) and return; `wait_host(slot)` delivers the results. While one
        slot's results drain over PCIe the other slot's kernels run."""
        b = int(packed.shape[0])
        if b > self.max_batch or b <= 0:
            raise ValueError(f"batch {b} outside 1..{self.max_batch}")
        h_states, h_policy, h_value, h_flags = self._slot_buffers(slot)
        src = packed if isinstance(packed, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(packed))
        h_states[:b].copy_(src)
        _lib.check(
            self.lib.at_engine_submit_host(self.handle, precision, _ptr(h_states), b, _ptr(h_policy), _ptr(h_value), _ptr(h_flags), slot,
                                           _stream_ptr(self.device)),
            "at_engine_submit_host",
        )
        self._slot_batch[slot] = b

    def wait_host(self, slot: int):
        """(probs ndarray [B,A], value ndarray [B], flags int) of the batch submitted on `slot`; views of that slot's pinned
        staging, valid until the slot is submitted to again."""
        _lib.check(self.lib.at_engine_wait_host(self.handle, slot), "at_engine_wait_host")
        _h_states, h_policy, h_value, h_flags = self._slot_buffers(slot)
        b = self._slot_batch[slot]
        return h_policy[:b].numpy(), h_value[:b].numpy(), int(h_flags[0].item())

    def set_kernel_timing(self, on: bool) -> None:
        """Record event pairs around the tower / FC kernels of the evaluations that follow (off by default)."""
        _lib.check(self.lib.at_engine_set_kernel_timing(self.handle, 1 if on else 0), "at_engine_set_kernel_timing")

    def kernel_times(self) -> dict | None:
        """Device milliseconds of the tower and FC kernels of the last tensor-core evaluation (waits for it); needs
        set_kernel_timing(True) before that evaluation."""
        t, f = C.c_float(), C.c_float()
        rc = self.lib.at_engine_kernel_ms(self.handle, C.byref(t), C.byref(f))
        if rc != _lib.AT_OK:
            return None
        return {"tower_ms": float(t.value), "fc_ms": float(f.value)}

    def bf16_trunk(self, packed: torch.Tensor) -> torch.Tensor:
        """Test hook: first shared-FC activation [B, N] computed by the tcgen05 kernels."""
        n = int(self.desc.shared_dims[0])
        out = torch.empty((packed.shape[0], n), dtype=torch.float32, device=self.device)
        _lib.check(
            self.lib.at_engine_bf16_trunk(self.handle, _ptr(packed), packed.shape[0], _ptr(out), _stream_ptr(self.device)),
            "at_engine_bf16_trunk",
        )
        return out

    def extract_features(self, packed: torch.Tensor, grid_channels: int, other_dim: int):
        """PackedState[B] on the device -> (grid [B,Cg,R,C], other [B,D]) on the device."""
        b = packed.shape[0]
        env = self.desc.env
        grid = torch.empty((b, grid_channels, env.rows, env.cols), dtype=torch.float32, device=self.device)
        other = torch.empty((b, other_dim), dtype=torch.float32, device=self.device)
        if b:
            _lib.check(
                self.lib.at_extract_features(
                    C.byref(self.desc.env), grid_channels, other_dim, _ptr(packed), b, _ptr(grid), _ptr(other),
                    _stream_ptr(self.device),
                ),
                "at_extract_features",
            )
        return grid, other
