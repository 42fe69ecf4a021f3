"""Pins the tcgen05 shared-memory descriptor conventions the convolution tower relies on:
no-swizzle K-major operands made of 8x16-byte core matrices, LBO = byte step between core
matrices along K, SBO = byte step between 8-row groups, and -- because rows are 16 bytes
apart when SBO = 128 -- an A operand that may start at ANY row of a taller image (the 3x3
taps of the convolution are row-shifted views of one padded board image). Exact integer data."""

import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tools"))

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [128, 64])
@pytest.mark.parametrize("shift", [0, 1, 16, 17, 34])
def test_row_shifted_interleaved_operand(n, shift):
    from probe_explore import image_interleaved, run

    rng = np.random.default_rng(shift * 7 + n)
    rows_a, k = 162, 64
    A = rng.integers(-2, 3, size=(rows_a, k)).astype(np.float32)
    B = rng.integers(-2, 3, size=(n, k)).astype(np.float32)
    out, status = run(
        image_interleaved(A), image_interleaved(B), n, k,
        dict(start=shift * 16, lbo=rows_a * 16, sbo=128, kstep=2 * rows_a * 16),
        dict(start=0, lbo=n * 16, sbo=128, kstep=2 * n * 16), 0,
    )
    assert status == 0
    assert np.array_equal(out, A[shift : shift + 128] @ B.T)


def test_mma_rate_probe_reports_the_pipe_rate():
    """Two issuing threads on different accumulators retire a 128x128x16 MMA about every 64 cycles."""
    import torch

    from alphatriangle_b200 import _lib

    lib = _lib.load_probe()
    src = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda")
    out = torch.zeros(8, dtype=torch.float32, device="cuda")
    assert lib.at_probe_mma_rate(128, 1024, 2, 0, 0, src.data_ptr(), 8, out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    per_issuer = out.cpu().numpy()
    assert (per_issuer > 100).all() and (per_issuer < 200).all(), per_issuer  # 128 when the pipe is saturated
    assert lib.at_probe_mma_rate(100, 1024, 2, 0, 0, src.data_ptr(), 8, out.data_ptr(), None) != 0  # n must be a multiple of 16
