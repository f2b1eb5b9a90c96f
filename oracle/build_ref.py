"""Recipe: compile the reference's own `extensions/hga` Cython module into oracle/_ref/ (test infrastructure).

The reference source is read where it lies (/root/reference/extensions/hga/hga_optimized.pyx); nothing is copied
into the repository.  Outputs (generated C + the extension module) go only into oracle/_ref/, which is git-ignored
but travels to the GPU box with the snapshot.  The GPU box has no /root/reference, so this script is a no-op there
(the prebuilt module is used).

Usage:  python oracle/build_ref.py            (idempotent; rebuilds only when the .pyx is newer)
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from pathlib import Path

HERE = Path(__file__).resolve().parent
OUT = HERE / "_ref"
REF_PYX = Path(os.environ.get("HGA_REFERENCE_ROOT", "/root/reference")) / "extensions" / "hga" / "hga_optimized.pyx"


def ref_module_path() -> Path:
    suffix = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
    return OUT / f"hga_optimized{suffix}"


def build(verbose: bool = False) -> Path | None:
    """Build oracle/_ref/hga_optimized*.so from the reference .pyx.  Returns the path, or None if unavailable."""
    so = ref_module_path()
    if not REF_PYX.exists():
        return so if so.exists() else None
    if so.exists() and so.stat().st_mtime >= REF_PYX.stat().st_mtime:
        return so
    import numpy

    OUT.mkdir(parents=True, exist_ok=True)
    c_file = OUT / "hga_optimized.c"
    run = lambda cmd: subprocess.run(cmd, check=True, stdout=None if verbose else subprocess.DEVNULL,
                                     stderr=None if verbose else subprocess.DEVNULL)
    run([sys.executable, "-m", "cython", "-3", str(REF_PYX), "-o", str(c_file)])
    # Same flags distutils would use for the reference's setup.py (no -ffast-math, no -march: plain x86-64 SSE2,
    # therefore no FMA contraction in the power sum).
    run(["gcc", "-O2", "-fPIC", "-shared", "-fwrapv", "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION",
         f"-I{sysconfig.get_paths()['include']}", f"-I{numpy.get_include()}", str(c_file), "-o", str(so), "-lm"])
    return so


def load():
    """Import the compiled reference module (or return None when it could not be built / is absent)."""
    so = build()
    if so is None or not so.exists():
        return None
    import importlib.util

    spec = importlib.util.spec_from_file_location("hga_optimized", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    p = build(verbose=True)
    print("reference extension:", p if p else "unavailable (no /root/reference and no prebuilt module)")
