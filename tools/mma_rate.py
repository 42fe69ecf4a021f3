"""What one B200 SM does with the tower's MMA shape (at_probe_mma_rate): issue rate of 128 x N x 16 bf16 MMAs with both
operands in shared memory, the cost of the per-slab wait / commit pattern, of concurrent traffic, and the hand-off
latencies between warps. Every variant runs on all 148 SMs at once; columns: mean / min / max over the CTAs.

    gpurun --timeout 600 -- 'python tools/mma_rate.py | tee gpurun_out/mma_rate.log'
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from alphatriangle_b200 import _lib  # noqa: E402

lib = _lib.load_probe()
src = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda")
grid = 148
out = torch.zeros(grid, dtype=torch.float32, device="cuda")


def run(n, mmas, issuers, copy, epi):
    for _ in range(2):
        rc = lib.at_probe_mma_rate(n, mmas, issuers, copy, epi, src.data_ptr(), grid, out.data_ptr(), None)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
    o = out.cpu()
    return f"{o.mean():7.1f} {o.min():7.1f} {o.max():7.1f}"


print("A. back-to-back MMAs, one thread per issuer; cycles per MMA of issuer 0")
print("   (copy = a third warp streams 16 KB bulk copies into shared memory; epi = warps reading tensor memory + storing to shared)")
for n in (64, 128, 256):
    for issuers in (1, 2):
        for copy in (0, 1):
            for epi in (0, 16):
                print(f"n={n:3d} issuers={issuers} copy={copy} epi={epi:2d}: {run(n, 4096, issuers, copy, epi)}")

print("B. the tower's per-slab pattern: per four MMAs 1 = mbarrier try_wait, 2 = tcgen05.commit, 4 = fence::after, 8 = epi warps poll a barrier")
for issuers in (1, 2):
    for flags in (0, 1, 2, 3, 7):
        print(f"issuers={issuers} flags={flags}: {run(128, 4096, issuers, 32 + flags, 0)}")
for epi in (8, 16):
    print(f"issuers=2 flags=3 epi={epi} (tensor-memory reads): {run(128, 4096, 2, 32 + 3, epi)}")
    print(f"issuers=2 flags=11 epi={epi} (polling): {run(128, 4096, 2, 32 + 11, epi)}")

print("C. latencies, cycles per round (a ping-pong round holds two hand-offs)")
for name, v in (("mbarrier ping-pong between two warps, try_wait", 10), ("mbarrier ping-pong between two warps, test_wait", 11),
                ("tcgen05.commit -> mbarrier, empty pipe", 12), ("one MMA + tcgen05.commit -> mbarrier", 13)):
    print(f"{name:48s}: {run(128, 2000, 1, v, 0)}")

print("D. tensor-memory read bandwidth, bytes per cycle per SM (w warps issuing back-to-back tcgen05.ld + wait::ld)")
for name, v in (("32x32b.x32", 64), ("32x32b.x16", 65), ("16x256b.x8", 66), ("16x128b.x8", 67), ("32x32b.x64", 68)):
    for w in (1, 4, 8, 16):
        print(f"{name:12s} warps={w:2d}: {run(128, 2000, 1, v, w)}")
