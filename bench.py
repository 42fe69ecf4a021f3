#!/usr/bin/env python
"""Leaf-evaluation throughput benchmark (BASELINE.json: states/sec at batch 4096 per B200).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload conv_only|default_tf]

One "step" = one pass of the hot path over one batch of synthetic states on each GPU:
PackedState records resident in HBM -> features -> AlphaTriangleNet forward -> policy
probabilities + expected values resident in HBM. Prints ONE JSON line (rank 0).

  value        states/s, device-timed (CUDA events), max over ranks, whole job
  e2e          the same through the host-buffer C-ABI call (pinned H2D of the records, D2H of probs/values inside the timed region)
  roofline     the dominant kernel against the measured bf16 tensor peak (MEASURED_PEAKS.json)
  cpu_baseline the oracle's restatement of the reference CPU evaluate_batch, timed on this box's host cores (rank 0, N=1 only)

`--impl reference` times that CPU restatement alone (all host threads) and prints the same line shape.
"""

from __future__ import annotations

import argparse
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

METRIC = "leaf evals/sec (states/sec) at batch 4096"
UNIT = "states/s"
FLOP_PER_STATE = {"conv_only": 167_855_872, "default_tf": 245_516_032}  # 2 FLOP/MAC, no padding credited
# dram__bytes_read.sum + dram__bytes_write.sum of tower_kernel from the `ncu --set full` capture in profiles/ (per launch, batch 4096):
# 1.73 MB read + 72.0 MB written (the 126 MB of FC operand tiles it produces mostly stay in the 126 MB L2)
# dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, one launch at batch 4096, from the committed ncu --set full
# captures (profiles/r01_ncu_full_step_final_raw.csv, profiles/r01_ncu_full_tower_tf_raw.csv); other batch sizes: not captured
TOWER_DRAM_BYTES_PER_LAUNCH = {("conv_only", 4096): 1_825_792 + 72_273_664, ("default_tf", 4096): 2_452_224 + 75_853_568}
# dominant kernel's own algorithmic FLOPs: conv_body + res_body (69,120 + 4,423,680 + 17,694,720 + 4 x 35,389,440), plus the two
# encoder layers (2 x 38,830,080) when they are fused into it
TOWER_FLOP_PER_STATE = {"conv_only": 163_745_280, "default_tf": 163_745_280 + 77_660_160}
WORKLOADS = {
    "conv_only": "conv-only AlphaTriangleNet (USE_TRANSFORMER=False), 8x15x3 env, batch 4096 synthetic states per GPU, features+forward",
    "default_tf": "default AlphaTriangleNet (2 transformer layers), 8x15x3 env, synthetic states, features+forward",
}


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


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

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
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_rate(workload: str, sample_states: int, budget_s: float, warmup: int = 1, steps: int | None = None):
    """The oracle's evaluate_batch (numpy features + torch-CPU fp32 forward + dict building) on synthetic states."""
    import torch

    import alphatriangle_b200 as at
    from alphatriangle_b200.synth import synth_state_dict, synth_states
    from oracle import model_ref

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    env, mc = at.EnvConfig(), model_config(workload)
    states = synth_states(sample_states, env, seed=1235)
    sd = synth_state_dict(at.AlphaTriangleNet(mc, env).state_dict(), 1)
    for _ in range(warmup):
        model_ref.evaluate_batch(sd, mc, states)
    times = []
    t_begin = time.perf_counter()
    while True:
        t0 = time.perf_counter()
        res = model_ref.evaluate_batch(sd, mc, states)
        times.append(time.perf_counter() - t0)
        assert len(res) == sample_states
        if steps is not None:
            if len(times) >= steps:
                break
        elif time.perf_counter() - t_begin >= budget_s and len(times) >= 3:
            break
    total = sum(times)
    return {
        "value": sample_states * len(times) / total, "unit": UNIT, "cores": cores, "kind": "port",
        "sample": f"{len(times)} x evaluate_batch({sample_states} synthetic states), {workload}, torch {torch.__version__} "
                  f"CPU fp32 with {cores} threads, numpy features, dict results",
        "ms_per_step": 1e3 * total / len(times), "p50_ms": 1e3 * statistics.median(times), "steps": len(times),
    }


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample = args.ref_sample
    r = cpu_reference_rate(args.workload, sample, budget_s=0.0, warmup=max(1, args.warmup), steps=args.steps)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": r["steps"],
        "warmup": max(1, args.warmup), "ms_per_step": r["ms_per_step"], "p50_ms": r["p50_ms"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload], "step": f"evaluate_batch on a bounded sample of {sample} states"},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    import torch.distributed as dist

    import alphatriangle_b200 as at
    from alphatriangle_b200 import _lib
    from alphatriangle_b200.packing import pack_states
    from alphatriangle_b200.synth import synth_state_dict, synth_states

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    env, mc = at.EnvConfig(), model_config(args.workload)
    B = args.batch
    nn = at.NeuralNetwork(mc, env, at.TrainConfig(), dev, precision=args.precision, max_batch=B)
    prec_name = "bf16" if nn.precision == _lib.PREC_BF16 else "fp32"

    # weights: rank 0 owns them (the trainer's role); every other rank receives the fp32 image over NCCL
    weight_bcast_ms = None
    if rank == 0:
        nn.set_weights(synth_state_dict(nn.get_weights(), 1))
    if world > 1:
        from alphatriangle_b200.dist import broadcast_weights

        broadcast_weights(nn, src=0)                    # first call also builds the NCCL communicator
        weight_bcast_ms = broadcast_weights(nn, src=0)  # steady state: one broadcast of the fp32 image + fold / repack on every rank

    # this rank's shard of the concurrent games (partitioned by game index; no data-path collective)
    states = synth_states(min(B, 1024), env, seed=1235 + rank)
    base = pack_states(states, env)
    import numpy as np

    packed_h = np.tile(base, (-(-B // base.shape[0]), 1))[:B]
    packed = torch.from_numpy(packed_h).to(dev)
    probs = torch.empty((B, nn.action_dim), dtype=torch.float32, device=dev)
    value = torch.empty((B,), dtype=torch.float32, device=dev)
    flags = torch.zeros((4,), dtype=torch.int32, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step():
        nn.engine.evaluate_packed(packed, nn.precision, probs, value, flags)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    launches_per_step = nn.engine.last_launches()

    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    with ClockSampler(local) as clocks:
        barrier()
        t_wall0 = time.perf_counter()
        for i in range(args.steps):
            flush.fill_(i & 0xFF)  # evict L2 between timed steps (outside the event pair)
            starts[i].record()
            step()
            ends[i].record()
        barrier()
        t_wall = time.perf_counter() - t_wall0
        # keep sampling clocks over a loaded stretch even when K is tiny
        step_ms = [s.elapsed_time(e) for s, e in zip(starts, ends)]
        t_load = time.perf_counter()
        while time.perf_counter() - t_load < 0.7:
            step()
        torch.cuda.synchronize(dev)
    total_ms = sum(step_ms)
    # second pass, same conditions, reading the engine's own event pair around the dominant kernel after every step
    tower_ms, fc_ms = [], []
    if nn.precision == _lib.PREC_BF16:
        for i in range(args.steps):
            flush.fill_(i & 0xFF)
            step()
            kt = nn.engine.kernel_times()
            if kt:
                tower_ms.append(kt["tower_ms"])
                fc_ms.append(kt["fc_ms"])
    if int(flags[0].item()) != 0:
        raise SystemExit(f"non-finite flag {int(flags[0].item())} raised during the benchmark")

    # e2e: host records in pinned memory -> C ABI host entry point -> probs/values back in pinned memory
    # Two batches in flight (at_engine_submit_host / at_engine_wait_host): every step still copies its own records up and
    # its own probs / values / flags down inside the timed region; the PCIe transfer of one batch runs under the kernels
    # of the next. `sync_ms_per_step` is the same with one blocking at_engine_evaluate_host call per step.
    e2e_steps = max(4, min(args.steps, 20))
    for _ in range(2):
        nn.engine.evaluate_host(packed_h, nn.precision)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        p_h, v_h, f_h = nn.engine.evaluate_host(packed_h, nn.precision)
    torch.cuda.synchronize(dev)
    e2e_sync_s = time.perf_counter() - t0
    for i in range(2):
        nn.engine.submit_host(packed_h, nn.precision, i)
    for i in range(2):
        nn.engine.wait_host(i)
    barrier()
    t0 = time.perf_counter()
    e2e_flags = 0
    for i in range(e2e_steps):
        if i >= 2:
            p_h, v_h, f_h = nn.engine.wait_host(i & 1)
            e2e_flags |= f_h
        nn.engine.submit_host(packed_h, nn.precision, i & 1)
    for i in range(e2e_steps - 2, e2e_steps):
        p_h, v_h, f_h = nn.engine.wait_host(i & 1)
        e2e_flags |= f_h
    e2e_s = time.perf_counter() - t0
    if e2e_flags:
        raise SystemExit(f"non-finite flag {e2e_flags} raised during the end-to-end loop")

    tmax = torch.tensor([total_ms, e2e_s * 1e3 / e2e_steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    total_ms_max, e2e_ms_max = float(tmax[0].item()), float(tmax[1].item())

    if rank == 0:
        peaks = measured_peaks()
        flops = FLOP_PER_STATE[args.workload] * B
        step_ms_avg = total_ms_max / args.steps
        achieved = flops / (step_ms_avg * 1e-3) / 1e12
        peak = peaks["sustained"] if args.steps * step_ms_avg > 1000 else peaks["burst"]
        kern = None
        if tower_ms:
            # dominant kernel = the convolution tower; its own CUDA-event time, averaged over a second pass of K steps
            t_ms = sum(tower_ms) / len(tower_ms)
            achieved = TOWER_FLOP_PER_STATE[args.workload] * B / (t_ms * 1e-3) / 1e12
            kern = {"name": "tower_kernel (conv_body + res_body, tcgen05)" if args.workload == "conv_only"
                    else "tower_tf_kernel (conv_body + res_body + transformer_body, tcgen05)", "ms": t_ms, "fc_ms": sum(fc_ms) / len(fc_ms),
                    "share_of_step": t_ms / step_ms_avg}
        line = {
            "metric": METRIC, "value": B * world * args.steps / (total_ms_max * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": step_ms_avg,
            "p50_ms": statistics.median(step_ms), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": prec_name, "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload], "batch_per_gpu": B, "global_batch": B * world, "precision_path": prec_name,
                       "sharding": "games partitioned by index across ranks, no data-path collective; weights via one NCCL broadcast",
                       "l2": "flushed between timed steps (256 MiB write outside the event pair)",
                       "weights": "synthetic (seeded), BatchNorm statistics perturbed"},
            "clocks": clocks.summary(),
            "e2e": {"value": B * world / (e2e_ms_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": B * 80,
                    "d2h_bytes_per_step": B * (nn.action_dim * 4 + 4) + 4, "ms_per_step": e2e_ms_max,
                    "api": "at_engine_submit_host / at_engine_wait_host, pinned host buffers, two batches in flight",
                    "sync_ms_per_step": e2e_sync_s * 1e3 / e2e_steps},
            "gpu_launches": int(launches_per_step * args.steps),
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                         "traffic": TOWER_DRAM_BYTES_PER_LAUNCH.get((args.workload, B)) if kern else None, "kernel": (kern or {}).get("name", "whole step (all kernels of the path)"),
                         "kernel_ms": (kern or {}).get("ms"), "fc_kernel_ms": (kern or {}).get("fc_ms"),
                         "kernel_share_of_step": (kern or {}).get("share_of_step"),
                         "peak_source": f"{peaks['source']} bf16 dense, {'sustained' if peak == peaks['sustained'] else 'burst'}",
                         "flop_per_state": FLOP_PER_STATE[args.workload]},
            "wall_s": t_wall,
        }
        if weight_bcast_ms is not None:
            line["weight_broadcast_ms"] = weight_bcast_ms
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_reference_rate(args.workload, args.ref_sample, budget_s=args.cpu_budget)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
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
    ap.add_argument("--precision", choices=["auto", "fp32", "bf16"], default="auto")
    ap.add_argument("--ref-sample", type=int, default=256, help="states per CPU-reference step")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU baseline work in the default run")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
