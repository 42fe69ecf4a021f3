"""Feature kernel alone at HBM-sized batches: achieved GB/s (algorithmic bytes / CUDA-event time) vs the measured peak."""
import ctypes as C
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import alphatriangle_b200 as at  # noqa: E402
from alphatriangle_b200 import _lib  # noqa: E402
from alphatriangle_b200.packing import pack_states  # noqa: E402
from alphatriangle_b200.synth import synth_states  # noqa: E402

lib = _lib.load()
env = at.EnvConfig()
base = pack_states(synth_states(1024, env, seed=3), env)
peak = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"] if (ROOT / "MEASURED_PEAKS.json").exists() else 6650.0
envd = _lib.env_desc(env)
for B in (4096, 65536, 1 << 20, 1 << 22):
    packed = torch.from_numpy(np.tile(base, (B // 1024, 1))).cuda()
    grid = torch.empty((B, 1, 8, 15), dtype=torch.float32, device="cuda")
    other = torch.empty((B, 30), dtype=torch.float32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run():
        _lib.check(lib.at_extract_features(C.byref(envd), 1, 30, C.c_void_p(packed.data_ptr()), B, C.c_void_p(grid.data_ptr()),
                                           C.c_void_p(other.data_ptr()), st), "at_extract_features")

    for _ in range(3):
        run()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ms = min(ts)
    nbytes = B * (80 + 480 + 120)
    print(f"features B={B}: {ms*1e3:.1f} us, {B/ms*1e3/1e9:.3f} G states/s, {nbytes/ms/1e6:.0f} GB/s algorithmic = {nbytes/ms/1e6/peak:.3f} of measured HBM peak {peak:.0f} GB/s")
