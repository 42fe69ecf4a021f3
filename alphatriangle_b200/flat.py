"""One flat buffer behind every float tensor of the network.

The reference publishes weights as a dict of ~94 CPU tensors through `ray.put` and one `set_weights.remote` per worker
(alphatriangle/training/worker_manager.py:169-209), and the trainer's optimiser updates ~94 separate parameter tensors
(alphatriangle/rl/core/trainer.py:204-310). Here the module's parameters and float buffers are VIEWS into one contiguous
float32 buffer laid out exactly like the engine's weight image (at_model_describe order, 16-byte-aligned entries), so

  * an optimiser step, `load_state_dict` or any other in-place update writes the image directly -- there is nothing to
    pack before the engine folds it or NCCL sends it;
  * a publication is ONE collective over the whole blob: the image plus a small int64 tail carrying the trainer's
    global step and the BatchNorm `num_batches_tracked` counters (the only non-float entries of the state_dict).
"""

from __future__ import annotations

import torch

HEADER_FIXED = 2  # int64 words: [0] trainer step / version, [1] image floats (sanity word)


class FlatWeights:
    def __init__(self, module: torch.nn.Module, table: list[tuple[str, int, int]], image_floats: int, device: torch.device):
        self.module = module
        self.table = table
        self.image_floats = int(image_floats)
        self.device = torch.device(device)
        sd = module.state_dict(keep_vars=True)
        self.int_keys = [k for k, v in sd.items() if not v.dtype.is_floating_point]
        padded = (self.image_floats + 3) & ~3  # int64 tail starts 16-byte aligned
        self.header_words = HEADER_FIXED + len(self.int_keys)
        self.blob = torch.zeros(padded + 2 * self.header_words, dtype=torch.float32, device=self.device)
        self.image = self.blob[: self.image_floats]
        self.header = self.blob[padded:].view(torch.int64)
        self._tensors: list[torch.Tensor] = []
        self._ptrs: list[int] = []
        self.adopt()

    def adopt(self) -> None:
        """Move every float tensor of the module into the blob (values preserved) and rebind it as a view."""
        sd = self.module.state_dict(keep_vars=True)
        names = {k for k, _o, _n in self.table}
        missing = [k for k, v in sd.items() if v.dtype.is_floating_point and k not in names]
        if missing:
            raise KeyError(f"weight table lacks {missing[:3]}...")
        self._tensors, self._ptrs = [], []
        base = self.blob.data_ptr()
        with torch.no_grad():
            for key, off, numel in self.table:
                t = sd[key]
                if t.numel() != numel:
                    raise ValueError(f"{key}: table says {numel} elements, module has {tuple(t.shape)}")
                view = self.blob[off : off + numel].view(t.shape)
                if t.data_ptr() != base + 4 * off or t.device != self.device:
                    view.copy_(t.detach().to(device=self.device, dtype=torch.float32))
                    t.data = view
                self._tensors.append(t)
                self._ptrs.append(base + 4 * off)
        self._ints = [sd[k] for k in self.int_keys]

    def aliased(self) -> bool:
        """False after something replaced a parameter's storage (module.to(), .half(), a new Parameter)."""
        return all(t.data_ptr() == p for t, p in zip(self._tensors, self._ptrs))

    def stamp(self) -> tuple:
        """Changes whenever any tensor was written in place (version counters) or re-bound."""
        return tuple(t._version for t in self._tensors)

    # ---- header tail ------------------------------------------------------------------
    def write_header(self, version: int) -> None:
        """Sender side, before the collective: version word and the integer buffers, all device-side."""
        with torch.no_grad():  # two small copies instead of one per word: publish() runs on the evaluating thread
            self.header[:2].copy_(torch.tensor([int(version), self.image_floats], dtype=torch.int64), non_blocking=True)
            if self._ints:
                words = torch.stack([t.reshape(-1)[0] for t in self._ints])
                self.header[HEADER_FIXED : HEADER_FIXED + len(self._ints)].copy_(words.to(device=self.device, dtype=torch.int64))

    def apply_header(self) -> None:
        """Receiver side, after the collective: integer buffers out of the tail, device-side (no host sync)."""
        with torch.no_grad():
            if self._ints:
                words = self.header[HEADER_FIXED : HEADER_FIXED + len(self._ints)]
                dsts = [t.reshape(-1)[0] if t.dim() else t for t in self._ints]
                torch._foreach_copy_(dsts, list(words.to(dsts[0].device).unbind()))
