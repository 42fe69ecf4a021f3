"""Run on a B200: which TMEM layout does tcgen05.mma expect for an A operand in tensor memory?"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parent))
from probe_explore import bf16_bits, image_interleaved  # noqa: E402

from alphatriangle_b200 import _lib  # noqa: E402


def run_ts(A, B, pack_mode, a_col, cols_per_kstep):
    lib = _lib.load_probe()
    dev = torch.device("cuda")
    n, k = B.shape
    da = torch.from_numpy(bf16_bits(A).astype(np.int16)).to(dev)
    b_img = image_interleaved(B)
    db = torch.from_numpy(b_img.view(np.uint8).reshape(-1).copy()).to(dev)
    out = torch.zeros((128, n), dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    rc = lib.at_probe_umma_ts(C.c_void_p(da.data_ptr()), C.c_void_p(db.data_ptr()), db.numel(), n, k, pack_mode, a_col, cols_per_kstep,
                              0, n * 16, 128, 2 * n * 16, C.c_void_p(out.data_ptr()), C.c_void_p(status.data_ptr()), None)
    _lib.check_probe(rc, "at_probe_umma_ts")
    torch.cuda.synchronize()
    return out.cpu().numpy(), int(status.item())


def main():
    rng = np.random.default_rng(1)
    for n, k in ((32, 128), (128, 64)):
        A = rng.integers(-2, 3, size=(128, k)).astype(np.float32)
        B = rng.integers(-2, 3, size=(n, k)).astype(np.float32)
        ref = A @ B.T
        for pack_mode in (0, 1):
            for cols in (8, 16):
                out, st = run_ts(A, B, pack_mode, 256, cols)
                print(f"TS n={n} k={k} pack_mode={pack_mode} cols_per_kstep={cols} status={st} max_err={np.abs(out - ref).max():.1f}")


if __name__ == "__main__":
    main()
