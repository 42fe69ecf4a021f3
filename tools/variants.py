"""Run the tower timing tools against several builds of the library (A/B kernel experiments).

    python tools/variants.py [name ...]        # name = suffix of alphatriangle_b200/libat_b200_<name>.so; "base" = the product library

Each build is timed in its own process (AT_B200_LIB selects it): conv-only tower and fused tower + encoder at batch 4096,
best of several launches, plus a parity spot check of the fused kernel against the product library's output.
"""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
names = sys.argv[1:] or ["base"]
for name in names:
    lib = ROOT / "alphatriangle_b200" / ("libat_b200.so" if name == "base" else f"libat_b200_{name}.so")
    env = dict(os.environ, AT_B200_LIB=str(lib))
    for tool, arg in (("tower_time.py", ["4096", "6"]), ("tf_time.py", ["4096"])):
        out = subprocess.run([sys.executable, str(ROOT / "tools" / tool), *arg], env=env, capture_output=True, text=True)
        line = (out.stdout.strip().splitlines() or [out.stderr.strip()[-300:]])[-1]
        print(f"{name:12s} {line}", flush=True)
