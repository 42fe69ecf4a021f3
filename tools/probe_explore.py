"""Run on a B200 (gpurun): which shared-memory descriptor conventions does tcgen05.mma honour?

Prints one line per hypothesis with the max abs error against an exact integer reference.
Used once during bring-up; the conventions that hold are then asserted by tests/test_gpu_probe.py.
"""

import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from alphatriangle_b200 import _lib  # noqa: E402


def bf16_bits(a: np.ndarray) -> np.ndarray:
    """float32 array of small integers -> uint16 bf16 bit patterns (exact)."""
    return (np.ascontiguousarray(a, dtype=np.float32).view(np.uint32) >> 16).astype(np.uint16)


def image_interleaved(mat: np.ndarray) -> np.ndarray:
    """[rows, K] -> no-swizzle K-major image [K/8][rows][8] (16-byte core-matrix rows, rows contiguous)."""
    rows, k = mat.shape
    return np.ascontiguousarray(bf16_bits(mat).reshape(rows, k // 8, 8).transpose(1, 0, 2))


def image_sw128(mat: np.ndarray) -> np.ndarray:
    """[rows, 64] -> 128-byte rows, 16-byte chunk j of row r stored at chunk j ^ (r % 8)."""
    rows, k = mat.shape
    assert k == 64
    src = bf16_bits(mat).reshape(rows, 8, 8)
    out = np.empty_like(src)
    for r in range(rows):
        for j in range(8):
            out[r, j ^ (r % 8)] = src[r, j]
    return out


def run(a_img, b_img, n, k, a, b, layout):
    lib = _lib.load_probe()
    dev = torch.device("cuda")
    da = torch.from_numpy(a_img.view(np.uint8).reshape(-1).copy()).to(dev)
    db = torch.from_numpy(b_img.view(np.uint8).reshape(-1).copy()).to(dev)
    out = torch.zeros((128, n), dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    rc = lib.at_probe_umma(
        C.c_void_p(da.data_ptr()), da.numel(), C.c_void_p(db.data_ptr()), db.numel(), n, k,
        a["start"], a["lbo"], a["sbo"], a["kstep"], b["start"], b["lbo"], b["sbo"], b["kstep"], layout,
        C.c_void_p(out.data_ptr()), C.c_void_p(status.data_ptr()), None,
    )
    _lib.check_probe(rc, "at_probe_umma")
    torch.cuda.synchronize()
    return out.cpu().numpy(), int(status.item())


def main():
    rng = np.random.default_rng(0)
    rows_a, k, n = 162, 64, 128
    A = rng.integers(-2, 3, size=(rows_a, k)).astype(np.float32)
    B = rng.integers(-2, 3, size=(n, k)).astype(np.float32)
    print("device:", torch.cuda.get_device_name(0))
    # ---- no swizzle ----
    a_img, b_img = image_interleaved(A), image_interleaved(B)
    # (the swapped reading, lbo = M/N step and sbo = K step, was tried once on a B200: it walks outside the
    #  image and faults with an illegal address, so it is not probed again)
    for hyp in ("lbo=K,sbo=MN",):
        for shift in (0, 8, 16, 1, 17, 33):
            ref = A[shift : shift + 128] @ B.T
            if hyp == "lbo=K,sbo=MN":
                a = dict(start=shift * 16, lbo=rows_a * 16, sbo=128, kstep=2 * rows_a * 16)
                b = dict(start=0, lbo=n * 16, sbo=128, kstep=2 * n * 16)
            else:
                a = dict(start=shift * 16, lbo=128, sbo=rows_a * 16, kstep=2 * rows_a * 16)
                b = dict(start=0, lbo=128, sbo=n * 16, kstep=2 * n * 16)
            out, st = run(a_img, b_img, n, k, a, b, 0)
            print(f"NOSWZ {hyp:14s} shift={shift:2d} status={st} max_err={np.abs(out - ref).max():.1f}")
    # ---- N = 64 (conv 32->64) ----
    B64 = B[:64]
    out, st = run(a_img, image_interleaved(B64), 64, k,
                  dict(start=17 * 16, lbo=rows_a * 16, sbo=128, kstep=2 * rows_a * 16),
                  dict(start=0, lbo=64 * 16, sbo=128, kstep=2 * 64 * 16), 0)
    print(f"NOSWZ N=64 shift=17 status={st} max_err={np.abs(out - A[17:145] @ B64.T).max():.1f}")
    # ---- 128-byte swizzle ----
    A2 = rng.integers(-2, 3, size=(144, 64)).astype(np.float32)
    a_sw, b_sw = image_sw128(A2), image_sw128(B)
    for shift in (0, 8, 1, 3):
        ref = A2[shift : shift + 128] @ B.T
        a = dict(start=shift * 128, lbo=16, sbo=1024, kstep=32)
        b = dict(start=0, lbo=16, sbo=1024, kstep=32)
        out, st = run(a_sw, b_sw, n, k, a, b, 2)
        print(f"SW128 shift={shift:2d} status={st} max_err={np.abs(out - ref).max():.1f}")


if __name__ == "__main__":
    main()
