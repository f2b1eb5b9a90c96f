"""In-tree build of libhga_b200.so (nvcc, sm_100a only).  Called by __graft_entry__.build() and, lazily, by _lib.load()
when the library is missing but nvcc is present.  The built .so is git-ignored and travels to the GPU box."""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIB = PKG / "libhga_b200.so"
SOURCES = [CSRC / "hga_api.cu"]
HEADERS = sorted(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "hga_b200.h"]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-shared",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def nvcc_path() -> str | None:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    return None


HASH = PKG / "libhga_b200.so.srchash"   # written next to the library: which sources and flags it was built from


def _source_hash() -> str:
    import hashlib
    h = hashlib.sha256(" ".join(NVCC_FLAGS).encode())
    for p in sorted(SOURCES + HEADERS):
        if p.exists():
            h.update(p.name.encode())
            h.update(p.read_bytes())
    return h.hexdigest()


def stale() -> bool:
    """The library is missing or was built from other sources.  Decided on content (a copy of the tree to another machine
    does not keep modification times in any useful order); modification times only when no hash was left by the build."""
    if not LIB.exists():
        return True
    if HASH.exists():
        return HASH.read_text().strip() != _source_hash()
    t = LIB.stat().st_mtime
    return any(p.stat().st_mtime > t for p in SOURCES + HEADERS if p.exists())


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not stale():
        return LIB
    nvcc = nvcc_path()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build libhga_b200.so")
    cmd = [nvcc, *NVCC_FLAGS, "-o", str(LIB), *map(str, SOURCES)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = PKG.parent / "gpurun_out"   # scratch, git-ignored: register / spill report of the last build
    log.mkdir(exist_ok=True)
    (log / "ptxas.log").write_text(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")
    HASH.write_text(_source_hash() + "\n")
    if verbose:
        print(res.stderr)
    return LIB
