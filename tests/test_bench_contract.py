"""bench.py's reference arm runs on CPU only: check the JSON line the driver parses."""

import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def test_reference_arm_prints_one_contract_line():
    out = subprocess.run(
        [sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2", "--warmup", "1", "--ref-sample", "8"],
        capture_output=True, text=True, timeout=300, cwd=ROOT,
    )
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "states/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["steps"] == 2 and d["n_gpus"] == 1
    # "reference" = the unmodified reference files under oracle/_ref (made by oracle/make_ref.py where /root/reference is mounted),
    # "port" = the oracle's restatement, used only when those files are absent
    from oracle import make_ref

    assert d["cpu_baseline"]["kind"] == ("reference" if make_ref.verify() else "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["metric"].endswith("at batch 4096") and d["config"]["batch_per_gpu"] == 4096
    assert d["e2e"] == {"value": d["value"], "unit": "states/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_exit_quietly():
    import os

    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_bench_line_schema_is_pinned():
    """Keys the driver and the judge read from the GPU arm's line: present in the code that builds it."""
    src = (ROOT / "bench.py").read_text()
    for key in ('"workloads"', '"sustained"', '"roofline"', '"cpu_baseline"', '"e2e"', '"gpu_launches"', '"clocks"', '"api"',
                '"traffic"', '"publication"', '"weight_broadcast_overlapped_ms"', '"step_time_inflation"', '"frac_of_sustained_peak"'):
        assert key in src, key
    import bench

    assert bench.metric_name(16384).endswith("at batch 16384")
    traffic, why = bench.dram_traffic_from_profiles("conv_only", 12345)
    assert traffic is None and "no ncu capture" in why
