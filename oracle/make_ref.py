"""Recipe for oracle/_ref/: the reference's OWN files for the leaf-evaluation path, unmodified (TEST INFRASTRUCTURE).

    python oracle/make_ref.py            # needs /root/reference; __graft_entry__.build() runs it when that tree is mounted

/root/reference is pure Python, so "building" its path means taking the four sub-packages the path imports
(alphatriangle/{config,features,nn,utils} plus the package __init__) byte for byte into oracle/_ref/alphatriangle/.
oracle/_ref/ is git-ignored -- no reference source enters the history -- but it is not gpurun-ignored, so it travels
to the GPU box like the built .so files do, and `bench.py --impl reference` and the `cpu_baseline` leg time THAT code
(kind "reference") with oracle/refstubs standing in for the two compiled third-party packages it imports for typing
(trianglengin, trimcts). A manifest with the SHA-256 of every copied file is written next to it; bench.py re-hashes the
copies before use and falls back to the restatement in oracle/model_ref.py (kind "port", labelled as such) if anything
is missing or differs.
"""

from __future__ import annotations

import hashlib
import json
import shutil
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = Path("/root/reference/alphatriangle")
DST = HERE / "_ref" / "alphatriangle"
SUBPACKAGES = ("config", "features", "nn", "utils")


def sha256(path: Path) -> str:
    return hashlib.sha256(path.read_bytes()).hexdigest()


def make() -> dict:
    if not SRC.is_dir():
        raise FileNotFoundError(f"{SRC} is not mounted here; oracle/_ref can only be made where the reference tree is")
    if DST.parent.exists():
        shutil.rmtree(DST.parent)
    DST.mkdir(parents=True)
    manifest = {}
    files = [SRC / "__init__.py"] + [p for sub in SUBPACKAGES for p in sorted((SRC / sub).glob("*.py"))]
    for src in files:
        rel = src.relative_to(SRC)
        out = DST / rel
        out.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(src, out)
        manifest[str(rel)] = sha256(out)
    (DST.parent / "MANIFEST.json").write_text(json.dumps({"source": str(SRC), "files": manifest}, indent=1))
    return manifest


def verify() -> bool:
    """True if oracle/_ref holds every file of the manifest with the recorded hash."""
    mf = DST.parent / "MANIFEST.json"
    if not mf.exists():
        return False
    try:
        files = json.loads(mf.read_text())["files"]
        return bool(files) and all((DST / rel).exists() and sha256(DST / rel) == h for rel, h in files.items())
    except Exception:
        return False


if __name__ == "__main__":
    m = make()
    print(f"oracle/_ref: {len(m)} files copied from {SRC}", file=sys.stderr)
