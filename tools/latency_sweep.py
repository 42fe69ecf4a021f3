"""p50 latency of one evaluate call against batch size (BASELINE.json's metric names "p50 batch latency"): device time of
at_engine_evaluate_packed (CUDA events) and wall time of the blocking host-buffer call, bf16 path, both workloads.

    gpurun --timeout 900 -- 'python tools/latency_sweep.py | tee gpurun_out/latency_sweep.jsonl'
"""
import json
import statistics
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import alphatriangle_b200 as at  # noqa: E402
from alphatriangle_b200.packing import pack_states  # noqa: E402
from alphatriangle_b200.synth import synth_state_dict, synth_states  # noqa: E402

env = at.EnvConfig()
states = synth_states(256, env, seed=5)
base = pack_states(states, env)
for workload in ("conv_only", "default_tf"):
    mc = at.ModelConfig(USE_TRANSFORMER=(workload == "default_tf"))
    nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision="auto", max_batch=16384)
    nn.set_weights(synth_state_dict(nn.get_weights(), 1))
    nn._sync_engine()
    from alphatriangle_b200 import _lib  # noqa: E402

    prec = _lib.PRECISION_NAMES[nn.precision]
    for B in (1, 8, 32, 128, 296, 512, 1024, 2048, 4096, 8192, 16384):
        packed_h = np.tile(base, (-(-B // 256), 1))[:B]
        packed_d = torch.from_numpy(packed_h).cuda()
        reps = 30
        for _ in range(5):
            nn.engine.evaluate_packed(packed_d, nn.precision)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
        for s, e in ev:
            s.record()
            nn.engine.evaluate_packed(packed_d, nn.precision)
            e.record()
        torch.cuda.synchronize()
        dev_ms = [s.elapsed_time(e) for s, e in ev]
        host_ms = []
        for _ in range(3):
            nn.engine.evaluate_host(packed_h, nn.precision)
        for _ in range(reps):
            t0 = time.perf_counter()
            nn.engine.evaluate_host(packed_h, nn.precision)
            host_ms.append((time.perf_counter() - t0) * 1e3)
        row = {"workload": workload, "precision": prec, "batch": B, "device_p50_ms": statistics.median(dev_ms),
               "host_call_p50_ms": statistics.median(host_ms), "states_per_s_device": B / (statistics.median(dev_ms) * 1e-3)}
        if B <= 128:  # what trimcts calls: evaluate_batch on state objects (packing, version check, kernels, result dicts)
            sts = (states * (-(-B // len(states))))[:B]
            api_ms = []
            for _ in range(3):
                nn.evaluate_batch(sts)
            for _ in range(reps):
                t0 = time.perf_counter()
                nn.evaluate_batch(sts)
                api_ms.append((time.perf_counter() - t0) * 1e3)
            row["evaluate_batch_p50_ms"] = statistics.median(api_ms)
        print(json.dumps(row), flush=True)
    del nn
