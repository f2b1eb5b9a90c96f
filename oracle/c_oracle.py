"""ctypes front end of oracle/hga_oracle.c — TEST INFRASTRUCTURE ONLY (see the header of hga_oracle.c).

`build()` compiles the C restatement with gcc (-O2 -ffp-contract=off: no FMA contraction, so the arithmetic is the
reference's) into oracle/_build/libhga_oracle.so; `lib()` loads it.
"""
from __future__ import annotations

import ctypes
import subprocess
from pathlib import Path
from typing import Optional, Tuple

import numpy as np

HERE = Path(__file__).resolve().parent
SRC = HERE / "hga_oracle.c"
OUT = HERE / "_build" / "libhga_oracle.so"

_lib = None


def build(force: bool = False) -> Path:
    if OUT.exists() and not force and OUT.stat().st_mtime >= SRC.stat().st_mtime:
        return OUT
    OUT.parent.mkdir(parents=True, exist_ok=True)
    subprocess.run(["gcc", "-O3", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", str(SRC), "-o", str(OUT), "-lm"],
                   check=True)
    return OUT


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(str(build()))
        dp, lp = ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_long)
        _lib.hga_oracle_max_threads.restype = ctypes.c_int
        _lib.hga_oracle_sosfilt.argtypes = [dp, ctypes.c_long, ctypes.c_long, ctypes.c_long, dp, ctypes.c_int, dp, dp,
                                            ctypes.c_long, ctypes.c_int]
        _lib.hga_oracle_sosfilt.restype = None
        _lib.hga_oracle_logpower.argtypes = [dp, ctypes.c_long, ctypes.c_long, lp, lp, ctypes.c_long, ctypes.c_double,
                                             dp, dp, dp, ctypes.c_int]
        _lib.hga_oracle_logpower.restype = None
        _lib.hga_oracle_extract_fused.argtypes = [dp, ctypes.c_long, ctypes.c_long, ctypes.c_long, dp, ctypes.c_int, dp,
                                                  dp, ctypes.c_int, dp, ctypes.c_long, ctypes.c_long, ctypes.c_double,
                                                  dp, dp, dp, ctypes.c_int]
        _lib.hga_oracle_extract_fused.restype = ctypes.c_long
    return _lib


def _p(a: Optional[np.ndarray], t=ctypes.c_double):
    return None if a is None else a.ctypes.data_as(ctypes.POINTER(t))


def max_threads() -> int:
    return int(lib().hga_oracle_max_threads())


def sosfilt(sos: np.ndarray, x: np.ndarray, zi: np.ndarray, n_threads: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """== scipy.signal.sosfilt(sos, x, axis=0, zi=zi) for x (T, C), zi (S, 2, C)."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    sos = np.ascontiguousarray(sos, dtype=np.float64)
    state = np.array(zi, dtype=np.float64, order="C", copy=True)
    y = np.empty_like(x)
    T, C = x.shape
    lib().hga_oracle_sosfilt(_p(x), T, C, C, _p(sos), sos.shape[0], _p(state), _p(y), C, n_threads)
    return y, state


def logpower(data: np.ndarray, start: np.ndarray, stop: np.ndarray, eps: float = 0.01, mu=None, sigma=None,
             n_threads: int = 0) -> np.ndarray:
    data = np.ascontiguousarray(data, dtype=np.float64)
    start = np.ascontiguousarray(start, dtype=np.int64)
    stop = np.ascontiguousarray(stop, dtype=np.int64)
    mu = None if mu is None else np.ascontiguousarray(np.broadcast_to(np.asarray(mu, np.float64).ravel(), (data.shape[1],)))
    sigma = None if sigma is None else np.ascontiguousarray(np.broadcast_to(np.asarray(sigma, np.float64).ravel(), (data.shape[1],)))
    out = np.empty((len(start), data.shape[1]), dtype=np.float64)
    lib().hga_oracle_logpower(_p(data), data.shape[1], data.shape[1], _p(start, ctypes.c_long), _p(stop, ctypes.c_long),
                              len(start), eps, _p(mu), _p(sigma), _p(out), n_threads)
    return out


def extract_fused(x: np.ndarray, sos_a: np.ndarray, zi_a: np.ndarray, sos_b: np.ndarray, zi_b: np.ndarray, L: int,
                  hop: int, eps: float = 0.01, mu=None, sigma=None, n_threads: int = 0) -> np.ndarray:
    """Fresh-extractor offline features of one stream, x (T, C) with T >= L. zi_*: untiled (S, 2)."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    T, C = x.shape
    F = (T - L) // hop + 1 if T >= L else 0
    out = np.empty((max(F, 0), C), dtype=np.float64)
    sos_a, sos_b = (np.ascontiguousarray(a, dtype=np.float64) for a in (sos_a, sos_b))
    zi_a, zi_b = (np.ascontiguousarray(a, dtype=np.float64) for a in (zi_a, zi_b))
    mu = None if mu is None else np.ascontiguousarray(np.asarray(mu, np.float64).ravel())
    sigma = None if sigma is None else np.ascontiguousarray(np.asarray(sigma, np.float64).ravel())
    got = lib().hga_oracle_extract_fused(_p(x), T, C, C, _p(sos_a), sos_a.shape[0], _p(zi_a), _p(sos_b), sos_b.shape[0],
                                         _p(zi_b), L, hop, eps, _p(mu), _p(sigma), _p(out), n_threads)
    assert got == out.shape[0]
    return out
