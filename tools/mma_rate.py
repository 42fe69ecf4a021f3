"""Measured tcgen05.mma issue rate of one SM under the tower's operand layout (at_probe_mma_rate)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from alphatriangle_b200 import _lib  # noqa: E402

lib = _lib.load()
src = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda")
grid = 148
out = torch.zeros(grid, dtype=torch.float32, device="cuda")
print("n issuers copy epi_warps -> cycles per MMA of issuer 0 (mean / min / max over 148 CTAs)")
for n in (64, 128, 256):
    for issuers in (1, 2):
        for copy in (0, 1):
            for epi in (0, 8, 16):
                for _ in range(2):
                    rc = lib.at_probe_mma_rate(n, 4096, issuers, copy, epi, src.data_ptr(), grid, out.data_ptr(), None)
                    assert rc == 0, _lib.last_error()
                    torch.cuda.synchronize()
                o = out.cpu()
                print(f"n={n:3d} issuers={issuers} copy={copy} epi={epi:2d}: {o.mean():7.1f} {o.min():7.1f} {o.max():7.1f}")

print("latency variants (cycles per round; ping-pong rounds hold two hand-offs)")
for name, v in (("mbarrier ping-pong, try_wait", 10), ("mbarrier ping-pong, test_wait", 11), ("commit -> barrier, empty pipe", 12), ("one MMA + commit -> barrier", 13)):
    for _ in range(2):
        rc = lib.at_probe_mma_rate(128, 2000, 1, v, 0, src.data_ptr(), grid, out.data_ptr(), None)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
    o = out.cpu()
    print(f"{name:34s}: {o.mean():7.1f} {o.min():7.1f} {o.max():7.1f}")
