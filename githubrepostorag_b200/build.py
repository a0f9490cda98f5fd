"""In-tree nvcc build of the C-ABI library (sm_100a only).

``python -m githubrepostorag_b200.build`` or ``__graft_entry__.build()``.  The ``.so`` lands in
``githubrepostorag_b200/_lib/`` (git-ignored, shipped to the GPU box with the snapshot).
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB_DIR = PKG / "_lib"
LIB_PATH = LIB_DIR / "libcoderag_b200.so"
STAMP = LIB_DIR / "libcoderag_b200.stamp"

NVCC_FLAGS = [
    "-O3", "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-shared",
]


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or "/usr/local/cuda/bin/nvcc"
    return cand if Path(cand).exists() else "nvcc"


def _sources_digest() -> str:
    h = hashlib.sha256()
    files = sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "coderag_b200.h"]
    for f in files:
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile ``csrc/coderag_b200.cu`` -> ``_lib/libcoderag_b200.so`` if sources changed."""
    LIB_DIR.mkdir(exist_ok=True)
    digest = _sources_digest()
    if not force and LIB_PATH.exists() and STAMP.exists() and STAMP.read_text().strip() == digest:
        return LIB_PATH
    cmd = [_nvcc(), *NVCC_FLAGS, "-lcuda", "-o", str(LIB_PATH), str(CSRC / "coderag_b200.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError(f"nvcc failed ({res.returncode}) building {LIB_PATH.name}")
    if verbose:
        print(res.stdout + res.stderr)
    STAMP.write_text(digest)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
