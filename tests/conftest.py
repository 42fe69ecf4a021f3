import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT, ROOT / "tests" / "golden"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_sessionstart(session):
    """A fresh checkout has no built artefacts (*.so is git-ignored): build them once, like __graft_entry__.build() does.
    nvcc cross-compiles without a GPU; on a box without the toolchain the prebuilt files that travelled with the tree are used."""
    import shutil
    import subprocess

    wanted = [ROOT / "alphatriangle_b200" / "libat_b200.so", ROOT / "alphatriangle_b200" / "libat_b200_probe.so",
              ROOT / "alphatriangle_b200" / "_pyhost.so", ROOT / "oracle" / "liboracle_features.so"]
    if all(w.exists() for w in wanted) or not shutil.which("make"):
        return
    jobs = str(min(8, os.cpu_count() or 1))
    if shutil.which("nvcc") or (ROOT / "alphatriangle_b200" / "libat_b200.so").exists():
        subprocess.run(["make", "-j", jobs], cwd=ROOT / "alphatriangle_b200" / "csrc", check=False)
    subprocess.run(["make", "-j", jobs], cwd=ROOT / "oracle", check=False)


GOLDEN_DIR = ROOT / "tests" / "golden"


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN_DIR
