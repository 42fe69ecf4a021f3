#!/usr/bin/env python
"""Leaf-evaluation throughput benchmark (BASELINE.json: states/sec at batch 4096 per B200).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload conv_only|default_tf] [--batch B]

One "step" = one pass of the hot path over one batch of synthetic states on each GPU:
PackedState records resident in HBM -> features -> AlphaTriangleNet forward -> policy
probabilities + expected values resident in HBM. Prints ONE JSON line (rank 0).

  value        states/s of the headline workload (BASELINE config 2: conv-only, 4096 states per GPU), device-timed with
               CUDA events, max over ranks, whole job
  e2e          the same through the host-buffer C-ABI call (pinned H2D of the records, D2H of probs/values in the timed region)
  roofline     the dominant kernel against the measured bf16 tensor peak (MEASURED_PEAKS.json)
  workloads    the other BASELINE configurations measured in the same run with the same method: the reference's default
               model (2 transformer layers) at 4096 and 16384 states (config 3) and, for N > 1, at 8192 states per GPU
               (config 4: 65,536 concurrent games on 8 GPUs) with a weight publication in flight
  sustained    >= 2 s of back-to-back steps of the headline workload (no L2 flush, no gaps), against the sustained peak
  api          NeuralNetwork.evaluate_batch(list of states): packing and result dicts included (what trimcts calls)
  cpu_baseline the reference's own evaluate_batch (oracle/_ref, unmodified; torch CPU + numba features) on this box's
               host cores (rank 0, N=1 only); falls back to the oracle's restatement, labelled "port"

`--impl reference` times that CPU code alone (all host threads) on the same workload and batch and prints the same
line shape.
"""

from __future__ import annotations

import argparse
import csv
import hashlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

UNIT = "states/s"
FLOP_PER_STATE = {"conv_only": 167_855_872, "default_tf": 245_516_032}  # 2 FLOP/MAC, no padding credited
# dominant kernel's own algorithmic FLOPs: conv_body + res_body (69,120 + 4,423,680 + 17,694,720 + 4 x 35,389,440), plus the two
# encoder layers (2 x 38,830,080) when they are fused into it
TOWER_FLOP_PER_STATE = {"conv_only": 163_745_280, "default_tf": 163_745_280 + 77_660_160}
TOWER_KERNEL = {"conv_only": "tower_kernel", "default_tf": "tower_tf_kernel"}
WORKLOADS = {
    "conv_only": "conv-only AlphaTriangleNet (USE_TRANSFORMER=False), 8x15x3 env, synthetic states, features+forward",
    "default_tf": "default AlphaTriangleNet (2 transformer layers), 8x15x3 env, synthetic states, features+forward",
}
KERNEL_SOURCES = ("alphatriangle_b200/csrc/at_bf16.cu", "alphatriangle_b200/csrc/at_bf16_tf.inc", "alphatriangle_b200/csrc/at_umma.cuh")


def metric_name(batch: int) -> str:
    return f"leaf evals/sec (states/sec) at batch {batch}"


def model_config(workload):
    import alphatriangle_b200 as at

    return at.ModelConfig(USE_TRANSFORMER=(workload == "default_tf"))


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"burst": float(d["bf16_tflops"]), "sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                "hbm": float(d["hbm_gbs"]), "source": "measured"}
    return {"burst": 1590.0, "sustained": 1400.0, "hbm": 6650.0, "source": "fallback"}


def kernel_source_hash() -> str:
    h = hashlib.sha256()
    for rel in KERNEL_SOURCES:
        h.update((ROOT / rel).read_bytes())
    return h.hexdigest()[:16]


def dram_traffic_from_profiles(workload: str, batch: int):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, one launch, read at run time from the committed
    `ncu --set full` capture profiles/r02_ncu_full_<workload>_b<batch>_raw.csv. The capture's side-car .json names the kernel
    sources' hash it was taken at: a capture of other code, another batch, or none at all gives null (plus the reason)."""
    raw = ROOT / "profiles" / f"r02_ncu_full_{workload}_b{batch}_raw.csv"
    meta = raw.with_suffix(".json")
    if not raw.exists() or not meta.exists():
        return None, "no ncu capture committed for this workload and batch"
    try:
        m = json.loads(meta.read_text())
        if m.get("kernel_source_hash") != kernel_source_hash():
            return None, f"capture is of kernel sources {m.get('kernel_source_hash')}, this build is {kernel_source_hash()}"
        rows = list(csv.reader(raw.open()))
        hdr, units = rows[0], rows[1]
        col = {h: i for i, h in enumerate(hdr)}
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        for r in rows[2:]:
            name = r[col["Kernel Name"]]
            if TOWER_KERNEL[workload] + "<" in name or name.endswith(TOWER_KERNEL[workload]):
                total = 0.0
                for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    total += float(r[col[key]].replace(",", "")) * scale[units[col[key]]]
                return int(total), f"profiles/{raw.name}"
        return None, f"{raw.name} holds no {TOWER_KERNEL[workload]} launch"
    except Exception as e:  # a malformed capture must not take the benchmark down
        return None, f"could not read {raw.name}: {e}"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, watts, reasons = [], [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
            try:
                watts.append(float(r[6]))
            except Exception:
                pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        out = {"sm_mhz": statistics.median(sm), "sm_min_mhz": min(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}
        if watts:
            out["power_w_median"] = statistics.median(watts)
        return out


# ---- CPU arm: the reference's own evaluate_batch ------------------------------------------------------------------

def _load_reference():
    """-> (kind, module namespace). kind "reference": the unmodified files under oracle/_ref (hash-checked against the
    manifest oracle/make_ref.py wrote); "port": the oracle's restatement."""
    from oracle import make_ref

    if make_ref.verify():
        for p in (str(ROOT / "oracle" / "refstubs"), str(ROOT / "oracle" / "_ref")):
            if p not in sys.path:
                sys.path.insert(0, p)
        os.environ.setdefault("NUMBA_CACHE_DIR", "/tmp/numba_cache_at_b200")
        try:
            from alphatriangle.config import ModelConfig as RefModelConfig, TrainConfig as RefTrainConfig
            from alphatriangle.features import extract_state_features as ref_extract
            from alphatriangle.nn import NeuralNetwork as RefNN
            from trianglengin import EnvConfig as RefEnvConfig

            return "reference", {"ModelConfig": RefModelConfig, "TrainConfig": RefTrainConfig, "extract": ref_extract, "NN": RefNN,
                                 "EnvConfig": RefEnvConfig}
        except Exception as e:  # e.g. numba missing on the box
            print(f"bench: oracle/_ref present but not importable ({e}); timing the port instead", file=sys.stderr)
    return "port", None


def cpu_reference_rate(workload: str, batch: int, budget_s: float, warmup: int = 1, steps: int | None = None, split: bool = True):
    """The reference's NeuralNetwork(device=cpu).evaluate_batch on `batch` synthetic states: numba features + torch-CPU fp32
    forward + dict building, all host threads. Timed per call; features / forward / post split measured on one extra pass."""
    import numpy as np
    import torch

    import alphatriangle_b200 as at
    from alphatriangle_b200.synth import synth_state_dict, synth_states

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    env = at.EnvConfig()
    ours_mc = model_config(workload)
    states = synth_states(batch, env, seed=1235)
    kind, ref = _load_reference()
    versions = {"torch": torch.__version__, "numpy": np.__version__}
    if kind == "reference":
        import numba

        versions["numba"] = numba.__version__
        mc = ref["ModelConfig"](USE_TRANSFORMER=(workload == "default_tf"))
        nn = ref["NN"](mc, ref["EnvConfig"](), ref["TrainConfig"](COMPILE_MODEL=False, DEVICE="cpu", WORKER_DEVICE="cpu"), torch.device("cpu"))
        nn.set_weights(synth_state_dict(nn.get_weights(), 1))
        run = lambda: nn.evaluate_batch(states)  # noqa: E731

        def split_once():
            t0 = time.perf_counter()
            feats = [ref["extract"](s, mc) for s in states]
            grid = torch.from_numpy(np.stack([f["grid"] for f in feats]))
            other = torch.from_numpy(np.stack([f["other_features"] for f in feats]))
            t1 = time.perf_counter()
            with torch.inference_mode():
                nn.model(grid, other)
            t2 = time.perf_counter()
            return t1 - t0, t2 - t1
    else:
        from oracle import model_ref

        sd = synth_state_dict(at.AlphaTriangleNet(ours_mc, env).state_dict(), 1)
        run = lambda: model_ref.evaluate_batch(sd, ours_mc, states)  # noqa: E731
        split_once = None
    import logging

    logging.disable(logging.WARNING)  # the reference logs per call; keep the JSON line the only output
    for _ in range(warmup):
        run()
    times = []
    t_begin = time.perf_counter()
    while True:
        t0 = time.perf_counter()
        res = run()
        times.append(time.perf_counter() - t0)
        assert len(res) == batch
        if steps is not None:
            if len(times) >= steps:
                break
        elif time.perf_counter() - t_begin >= budget_s and len(times) >= 3:
            break
    total = sum(times)
    out = {
        "value": batch * len(times) / total, "unit": UNIT, "cores": cores, "threads": torch.get_num_threads(), "kind": kind,
        "sample": f"{len(times)} x evaluate_batch({batch} synthetic states), {workload}, "
                  + ("unmodified reference (oracle/_ref: alphatriangle.nn.NeuralNetwork on cpu, numba features)" if kind == "reference"
                     else "oracle/model_ref.py restatement (numpy features)")
                  + f", torch {torch.__version__} CPU fp32 with {cores} threads, dict results",
        "batch": batch, "workload": workload, "versions": versions,
        "ms_per_step": 1e3 * total / len(times), "p50_ms": 1e3 * statistics.median(times), "best_ms": 1e3 * min(times), "steps": len(times),
    }
    if split and split_once is not None:
        f_s, fw_s = split_once()
        out["split_ms"] = {"features": 1e3 * f_s, "forward": 1e3 * fw_s, "post": max(0.0, out["p50_ms"] - 1e3 * (f_s + fw_s))}
    logging.disable(logging.NOTSET)
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    batch = args.ref_sample or args.batch
    r = cpu_reference_rate(args.workload, batch, budget_s=0.0, warmup=max(1, args.warmup), steps=args.steps, split=False)
    line = {
        "impl": "reference", "metric": metric_name(args.batch), "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"],
        "warmup": max(1, args.warmup), "ms_per_step": r["ms_per_step"], "p50_ms": r["p50_ms"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload], "batch_per_gpu": args.batch,
                   "step": f"one evaluate_batch call on {batch} states on the host cores (rank 0 only; other ranks idle)"},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "threads", "kind", "sample", "versions")},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---- GPU arm ----------------------------------------------------------------------------------------------------

class Workload:
    """One (model, batch) configuration on this rank's GPU: network, resident inputs / outputs, measurement methods."""

    def __init__(self, name, batch, precision, dev, rank, world, weights_from_rank0=True):
        import numpy as np
        import torch

        import alphatriangle_b200 as at
        from alphatriangle_b200 import _lib
        from alphatriangle_b200.packing import pack_states
        from alphatriangle_b200.synth import synth_state_dict, synth_states

        self.name, self.B, self.dev, self.rank, self.world = name, batch, dev, rank, world
        self.torch, self._lib = torch, _lib
        env, mc = at.EnvConfig(), model_config(name)
        self.nn = at.NeuralNetwork(mc, env, at.TrainConfig(), dev, precision=precision, max_batch=batch)
        self.prec_name = _lib.PRECISION_NAMES[self.nn.precision]
        # weights: rank 0 owns them (the trainer's role); every other rank receives the image over NCCL
        if rank == 0 or not weights_from_rank0:
            self.nn.set_weights(synth_state_dict(self.nn.get_weights(), 1))
        self.weight_bcast_ms = None
        if world > 1 and weights_from_rank0:
            from alphatriangle_b200.dist import broadcast_weights

            broadcast_weights(self.nn, src=0)                         # first call also builds the NCCL communicator
            self.weight_bcast_ms = broadcast_weights(self.nn, src=0)  # steady state, GPU otherwise idle: collective + fold / repack
        # this rank's shard of the concurrent games (partitioned by game index; no data-path collective)
        self.states = synth_states(min(batch, 1024), env, seed=1235 + rank)
        base = pack_states(self.states, env)
        self.packed_h = np.tile(base, (-(-batch // base.shape[0]), 1))[:batch]
        self.packed = torch.from_numpy(self.packed_h).to(dev)
        self.probs = torch.empty((batch, self.nn.action_dim), dtype=torch.float32, device=dev)
        self.value = torch.empty((batch,), dtype=torch.float32, device=dev)
        self.flags = torch.zeros((4,), dtype=torch.int32, device=dev)

    def step(self):
        self.nn.engine.evaluate_packed(self.packed, self.nn.precision, self.probs, self.value, self.flags)

    def barrier(self):
        if self.world > 1:
            self.torch.distributed.barrier()
        self.torch.cuda.synchronize(self.dev)

    def check_flags(self, where):
        f = int(self.flags[0].item())
        if f:
            raise SystemExit(f"non-finite / pipeline flag {f:#x} raised during {where} ({self.name}@{self.B})")

    def device_timed(self, steps, warmup, flush, clocks=None):
        """K steps, each bracketed by its own CUDA event pair on the launching stream, L2 flushed between steps."""
        torch = self.torch
        for _ in range(max(3, warmup)):
            self.step()
        self.barrier()
        self.launches_per_step = self.nn.engine.last_launches()
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        self.barrier()
        t0 = time.perf_counter()
        for i in range(steps):
            flush.fill_(i & 0xFF)  # evict L2 between timed steps (outside the event pair)
            starts[i].record()
            self.step()
            ends[i].record()
        self.barrier()
        wall = time.perf_counter() - t0
        ms = [s.elapsed_time(e) for s, e in zip(starts, ends)]
        self.check_flags("the timed steps")
        return ms, wall

    def kernel_times(self, steps, flush):
        """Second pass, same conditions: the engine's own event pair around the dominant kernel, read after every step."""
        tower, fc = [], []
        if self.nn.precision not in self._lib.TENSOR_CORE_PRECISIONS:
            return None
        self.nn.engine.set_kernel_timing(True)  # only for this pass: the timed steps above run without the extra events
        for i in range(steps):
            flush.fill_(i & 0xFF)
            self.step()
            kt = self.nn.engine.kernel_times()
            if kt:
                tower.append(kt["tower_ms"])
                fc.append(kt["fc_ms"])
        self.nn.engine.set_kernel_timing(False)
        if not tower:
            return None
        return {"tower_ms": sum(tower) / len(tower), "fc_ms": sum(fc) / len(fc)}

    def e2e(self, steps):
        """Host records in pinned memory -> C ABI -> probs / values back in pinned memory, two batches in flight
        (at_engine_submit_host / at_engine_wait_host): every step copies its own records up and its own results down inside
        the timed region. Also the same with one blocking at_engine_evaluate_host per step."""
        torch, eng, prec = self.torch, self.nn.engine, self.nn.precision
        for _ in range(2):
            eng.evaluate_host(self.packed_h, prec)
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            eng.evaluate_host(self.packed_h, prec)
        torch.cuda.synchronize(self.dev)
        sync_s = time.perf_counter() - t0
        for i in range(2):
            eng.submit_host(self.packed_h, prec, i)
        for i in range(2):
            eng.wait_host(i)
        self.barrier()
        t0 = time.perf_counter()
        flags = 0
        for i in range(steps):
            if i >= 2:
                flags |= eng.wait_host(i & 1)[2]
            eng.submit_host(self.packed_h, prec, i & 1)
        for i in range(steps - 2, steps):
            flags |= eng.wait_host(i & 1)[2]
        pipe_s = time.perf_counter() - t0
        if flags:
            raise SystemExit(f"non-finite flag {flags} raised during the end-to-end loop ({self.name}@{self.B})")
        return pipe_s * 1e3 / steps, sync_s * 1e3 / steps

    def sustained(self, seconds):
        """Back-to-back steps for `seconds` (no flush, no gaps): states/s over the whole stretch, device-timed."""
        torch = self.torch
        for _ in range(3):
            self.step()
        self.barrier()
        n_chunk = 200
        begin, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        done = 0
        begin.record()
        t0 = time.perf_counter()
        while True:
            for _ in range(n_chunk):
                self.step()
            done += n_chunk
            torch.cuda.current_stream(self.dev).synchronize()
            if time.perf_counter() - t0 >= seconds:
                break
        end.record()
        torch.cuda.synchronize(self.dev)
        self.check_flags("the sustained run")
        return done, begin.elapsed_time(end)

    def api_level(self, steps):
        """NeuralNetwork.evaluate_batch on Python state objects: host packing, H2D, kernels, D2H, result dicts."""
        n = self.B
        states = (self.states * (-(-n // len(self.states))))[:n]
        self.nn.evaluate_batch(states[:64])
        times = []
        for _ in range(steps):
            t0 = time.perf_counter()
            res = self.nn.evaluate_batch(states)
            times.append(time.perf_counter() - t0)
            assert len(res) == n
        return {"value": n / statistics.median(times), "unit": UNIT, "p50_ms": 1e3 * statistics.median(times), "batch": n,
                "api": "NeuralNetwork.evaluate_batch(list[state]) -> list[(dict, float)]: C packer, pinned H2D, kernels, D2H, C result builder"}


def reduce_max(torch, dev, world, values):
    t = torch.tensor(values, dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


def roofline_record(name, B, step_ms, kt, peaks, long_run=False):
    peak = peaks["sustained"] if long_run else peaks["burst"]
    if kt:
        achieved = TOWER_FLOP_PER_STATE[name] * B / (kt["tower_ms"] * 1e-3) / 1e12
        kernel = ("tower_kernel (conv_body + res_body, tcgen05)" if name == "conv_only"
                  else "tower_tf_kernel (conv_body + res_body + transformer_body, tcgen05)")
        traffic, traffic_src = dram_traffic_from_profiles(name, B)
    else:
        achieved = FLOP_PER_STATE[name] * B / (step_ms * 1e-3) / 1e12
        kernel, traffic, traffic_src = "whole step (all kernels of the path)", None, "no single dominant kernel timed"
    return {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
            "traffic_source": traffic_src, "kernel": kernel, "kernel_ms": kt["tower_ms"] if kt else None,
            "fc_kernel_ms": kt["fc_ms"] if kt else None, "kernel_share_of_step": kt["tower_ms"] / step_ms if kt else None,
            "peak_source": f"{peaks['source']} bf16 dense, {'sustained' if long_run else 'burst'}",
            "flop_per_state": FLOP_PER_STATE[name], "kernel_flop_per_state": TOWER_FLOP_PER_STATE[name] if kt else None}


def measure_secondary(name, batch, args, dev, rank, world, flush, peaks, with_publication=False):
    """One entry of `workloads`: same timing method as the headline, fewer steps."""
    import torch

    w = Workload(name, batch, args.precision, dev, rank, world)
    steps = max(5, min(args.steps, 20))
    ms, _wall = w.device_timed(steps, args.warmup, flush)
    kt = w.kernel_times(min(steps, 10), flush)
    e2e_ms, e2e_sync_ms = w.e2e(max(4, min(steps, 10)))
    rec_pub = None
    if with_publication and world > 1:
        rec_pub = measure_publication(w, steps, flush)
    total_ms, e2e_ms_max = reduce_max(torch, dev, world, [sum(ms), e2e_ms])
    step_ms = total_ms / steps
    rec = {"workload": WORKLOADS[name], "batch_per_gpu": batch, "global_batch": batch * world, "metric": metric_name(batch),
           "states_per_s": batch * world * steps / (total_ms * 1e-3), "ms_per_step": step_ms, "p50_ms": statistics.median(ms), "steps": steps,
           "dtype": w.prec_name, "gpu_launches_per_step": int(w.launches_per_step),
           "e2e": {"value": batch * world / (e2e_ms_max * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms_max, "sync_ms_per_step": e2e_sync_ms,
                   "h2d_bytes_per_step": batch * 80, "d2h_bytes_per_step": batch * (w.nn.action_dim * 4 + 4) + 4},
           "roofline": roofline_record(name, batch, step_ms, kt, peaks)}
    if w.weight_bcast_ms is not None:
        rec["weight_broadcast_ms"] = w.weight_bcast_ms
    if rec_pub:
        rec["publication"] = rec_pub
    del w
    torch.cuda.empty_cache()
    return rec


def measure_publication(w, steps, flush):
    """Weight publication with evaluation running: rank 0 publishes on a side stream while
    every rank keeps evaluating on its main stream. Reports the publication's own duration (side-stream events) and the
    evaluation's step time with and without publications in flight."""
    import torch

    from alphatriangle_b200.dist import Publisher

    pub = Publisher(w.nn)
    pub.publish(src=0)
    pub.wait()
    w.barrier()
    n, every, window = 160, 32, 8

    # ONE run, timed step by step (event to event, so a step's time includes the host gap in front of it -- publish() is
    # called from the evaluating thread). A publication's cost is what the `window` steps after it take beyond the median
    # of the other steps of the same 32-step stretch: a with/without comparison of two separate runs measures the power
    # cap's drift between them (step times move by 3 % over 100 steps) as much as the publication.
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    pubs = []
    w.barrier()
    evs[0].record()
    for i in range(n):
        if i % every == 2:  # every rank takes part in the collective; the main stream never waits for it
            p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            p0.record(pub.stream)
            pub.publish(src=0)
            p1.record(pub.stream)
            pubs.append((p0, p1))
        w.step()
        evs[i + 1].record()
        torch.cuda.current_stream(w.dev).synchronize()  # the host consumes each batch (never waits for the side stream)
    pub.wait()
    w.barrier()
    step_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(n)]
    pub_ms = [a.elapsed_time(b) for a, b in pubs]
    lost, quiet = [], []
    for k in range(0, n, every):
        rest = step_ms[k + 2 + window:k + every] + step_ms[k:k + 2]
        base = statistics.median(rest)
        quiet += rest
        lost.append(sum(v - base for v in step_ms[k + 2:k + 2 + window]))
    base_ms, with_ms, lost_ms = statistics.median(quiet), sum(step_ms) / n, sum(lost) / len(lost)
    w.check_flags("the publication run")
    # after the last publication every rank must hold rank 0's weights: the same probe batch gives the same bits everywhere
    from alphatriangle_b200.packing import pack_states
    from alphatriangle_b200.synth import synth_states

    probe = torch.from_numpy(pack_states(synth_states(256, w.nn.env_config, seed=4242), w.nn.env_config)).to(w.dev)
    pp, pv, _pf = w.nn.engine.evaluate_packed(probe, w.nn.precision)
    digest = torch.stack([pp.view(torch.int32).sum(dtype=torch.int64), pv.view(torch.int32).sum(dtype=torch.int64)])
    digests = [torch.zeros_like(digest) for _ in range(w.world)]
    torch.distributed.all_gather(digests, digest)
    identical = all(bool(torch.equal(d, digests[0])) for d in digests)
    vals = reduce_max(torch, w.dev, w.world, [base_ms, with_ms, statistics.median(pub_ms), lost_ms])
    return {"step_ms_alone": vals[0], "step_ms_with_publications": vals[1], "step_time_inflation": vals[3] / (every * vals[0]),
            "eval_ms_lost_per_publication": vals[3],
            "method": "one run, per-step device times; lost = what the 8 steps after a publication take beyond the median of the "
                      "other steps of the same 32-step stretch (host time of publish() on the evaluating thread included)",
            "weight_broadcast_overlapped_ms": vals[2], "publications": len(pub_ms), "every_n_steps": every,
            "replicas_bit_identical_after_publication": identical,
            "how": "rank 0's flat image + header broadcast (one NCCL collective) and fold / repack on a side stream of every rank "
                   "while the main stream evaluates; evaluations switch weight sets between calls (at_engine_publish_weights)"}


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = measured_peaks()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2
    B = args.batch

    # ---- headline workload --------------------------------------------------------------------------------------
    w = Workload(args.workload, B, args.precision, dev, rank, world)
    with ClockSampler(local) as clocks:
        step_ms, t_wall = w.device_timed(args.steps, args.warmup, flush)
        # keep sampling clocks over a loaded stretch even when K is tiny
        t_load = time.perf_counter()
        while time.perf_counter() - t_load < 0.7:
            w.step()
        torch.cuda.synchronize(dev)
    kt = w.kernel_times(args.steps, flush)
    e2e_steps = max(4, min(args.steps, 20))
    e2e_ms, e2e_sync_ms = w.e2e(e2e_steps)
    total_ms_max, e2e_ms_max = reduce_max(torch, dev, world, [sum(step_ms), e2e_ms])
    sustained = None
    if args.sustained > 0:
        with ClockSampler(local) as sclocks:
            done, sus_ms = w.sustained(args.sustained)
        (sus_ms_max,) = reduce_max(torch, dev, world, [sus_ms])
        rate = B * world * done / (sus_ms_max * 1e-3)
        sustained = {"seconds": sus_ms_max * 1e-3, "steps": done, "states_per_s": rate, "ms_per_step": sus_ms_max / done,
                     "tflops": FLOP_PER_STATE[args.workload] * rate / world / 1e12, "peak": peaks["sustained"],
                     "frac_of_sustained_peak": FLOP_PER_STATE[args.workload] * rate / world / 1e12 / peaks["sustained"],
                     "l2": "not flushed (steady state: weights L2-resident, inputs re-read)", "clocks": sclocks.summary()}
    api = w.api_level(3) if (world == 1 and not args.no_api) else None
    bcast_ms = w.weight_bcast_ms
    launches = int(w.launches_per_step)
    action_dim = w.nn.action_dim
    prec_name = w.prec_name
    del w
    torch.cuda.empty_cache()

    # ---- the other BASELINE configurations, same run, same method ------------------------------------------------
    workloads = {}
    if not args.no_workloads:
        plan = [("default_tf", 4096), ("default_tf", 16384)] if world == 1 else [("default_tf", 8192)]
        for name, batch in plan:
            if (name, batch) == (args.workload, B):
                continue
            workloads[f"{name}@{batch}"] = measure_secondary(name, batch, args, dev, rank, world, flush, peaks, with_publication=world > 1)

    if rank == 0:
        step_ms_avg = total_ms_max / args.steps
        line = {
            "metric": metric_name(B), "value": B * world * args.steps / (total_ms_max * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": step_ms_avg,
            "p50_ms": statistics.median(step_ms), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": prec_name, "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload], "batch_per_gpu": B, "global_batch": B * world, "precision_path": prec_name,
                       "sharding": "games partitioned by index across ranks, no data-path collective; weights via one NCCL broadcast",
                       "l2": "flushed between timed steps (256 MiB write outside the event pair)",
                       "weights": "synthetic (seeded), BatchNorm statistics perturbed"},
            "clocks": clocks.summary(),
            "e2e": {"value": B * world / (e2e_ms_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": B * 80,
                    "d2h_bytes_per_step": B * (action_dim * 4 + 4) + 4, "ms_per_step": e2e_ms_max,
                    "api": "at_engine_submit_host / at_engine_wait_host, pinned host buffers, two batches in flight",
                    "sync_ms_per_step": e2e_sync_ms},
            "gpu_launches": int(launches * args.steps),
            "roofline": roofline_record(args.workload, B, step_ms_avg, kt, peaks, long_run=args.steps * step_ms_avg > 1000),
            "wall_s": t_wall,
        }
        if sustained:
            line["sustained"] = sustained
        if workloads:
            line["workloads"] = workloads
        if api:
            line["api"] = api
        if bcast_ms is not None:
            line["weight_broadcast_ms"] = bcast_ms
        if world == 1 and not args.no_cpu_baseline:
            share = args.cpu_budget / 3.0
            r = cpu_reference_rate(args.workload, B, budget_s=share)
            line["cpu_baseline"] = {k: r[k] for k in r if k not in ("steps",)}
            rows = {}
            for name, batch in (("default_tf", 64), ("conv_only", 64)):  # BASELINE config 1 and its conv-only twin
                rr = cpu_reference_rate(name, batch, budget_s=share / 2)
                rows[f"{name}@{batch}"] = {k: rr[k] for k in ("value", "unit", "p50_ms", "best_ms", "kind", "cores", "threads", "split_ms")
                                           if k in rr}
            line["cpu_baseline"]["rows"] = rows
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=list(WORKLOADS), default="conv_only")
    ap.add_argument("--batch", type=int, default=4096, help="states per GPU per step")
    ap.add_argument("--precision", choices=["auto", "fp32", "bf16", "f16"], default="auto")
    ap.add_argument("--ref-sample", type=int, default=0, help="states per CPU-reference step (0 = --batch, like for like)")
    ap.add_argument("--cpu-budget", type=float, default=18.0, help="seconds of CPU baseline work in the default run")
    ap.add_argument("--sustained", type=float, default=2.5, help="seconds of back-to-back steps for the sustained record (0 = skip)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-workloads", action="store_true", help="headline workload only")
    ap.add_argument("--no-api", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
