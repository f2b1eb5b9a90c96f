"""Host-side filter design for the HGA cascade (coefficients are an INPUT of the kernels).

Mirrors `HighGammaExtractor.create_filter(...)["sos"]` (/root/reference/local/units.py:122-132,138-143).  The reference
designs through mne 1.6.1 `create_filter(method='iir', iir_params={'order': 8, 'ftype': 'butter'})`, whose IIR branch
reduces to `scipy.signal.iirfilter(N=order, Wn, btype, ftype='butter', output='sos')` with `l_freq > h_freq` meaning
band-stop.  If mne is importable it is used (exactly the reference's call); otherwise scipy is called directly.
"""
from __future__ import annotations

from functools import lru_cache
from typing import Tuple

import functools

import numpy as np
from scipy.signal import iirfilter, sosfilt_zi


@lru_cache(maxsize=64)
def _design(fs: float, l_freq: float, h_freq: float, order: int) -> Tuple[bytes, int]:
    try:  # the reference's own call, when its dependency is installed
        import mne  # type: ignore

        iir_params, method = mne.filter._check_method("iir", {"order": order, "ftype": "butter"})
        sos = mne.filter.create_filter(None, fs, l_freq, h_freq, "auto", "auto", "auto", method, iir_params, "zero",
                                       "hamming", "firwin", verbose=30)["sos"]
    except ImportError:
        nyq = fs / 2.0
        if l_freq < h_freq:
            sos = iirfilter(order, np.array([l_freq, h_freq]) / nyq, btype="bandpass", ftype="butter", output="sos")
        else:
            sos = iirfilter(order, np.array([h_freq, l_freq]) / nyq, btype="bandstop", ftype="butter", output="sos")
    sos = np.ascontiguousarray(sos, dtype=np.float64)
    return sos.tobytes(), sos.shape[0]


def create_filter_sos(fs: float, l_freq: float, h_freq: float, order: int = 8) -> np.ndarray:
    """(n_sections, 6) float64 SOS matrix; cached per (fs, l, h, order) — the reference re-designs per trial
    (prepare_corpus.py:48), which is pure overhead."""
    raw, n = _design(float(fs), float(l_freq), float(h_freq), int(order))
    return np.frombuffer(raw, dtype=np.float64).reshape(n, 6).copy()


def design_hga_filters(fs: float, l_freq: float = 70, h_freq: float = 170) -> Tuple[np.ndarray, np.ndarray]:
    """(hg_filter, fh_filter) of local/units.py:124-126: band-pass l..h Hz and the 118-122 Hz band-stop."""
    return create_filter_sos(fs, l_freq, h_freq), create_filter_sos(fs, 122, 118)


def initial_state(sos: np.ndarray) -> np.ndarray:
    """sosfilt_zi(sos): (n_sections, 2), NOT scaled by the first sample (local/units.py:128-129)."""
    return np.ascontiguousarray(sosfilt_zi(sos), dtype=np.float64)


def tiled_zi(sos: np.ndarray, nb_electrodes: int) -> np.ndarray:
    """The reference's (n_sections, 2, C) tiling (local/units.py:131-132)."""
    zi = initial_state(sos)
    return np.repeat(zi, nb_electrodes, axis=-1).reshape([zi.shape[0], zi.shape[1], -1])


def window_geometry(fs: int, window_length: float, window_shift: float) -> Tuple[int, int]:
    """(L, hop) in samples with the reference's C-float arithmetic (hga_optimized.pyx:76-77: `<int>(float * int)`)."""
    f32 = np.float32
    return int(f32(window_length) * f32(fs)), int(f32(window_shift) * f32(fs))


def window_indices(rows: int, sr: int, window_length: float, window_shift: float) -> Tuple[np.ndarray, np.ndarray]:
    """(start, stop) of every window exactly as compute_log_power_features derives them (hga_optimized.pyx:36,43-44):
    float32 products, libm floor/round on the promoted double.  Cached per geometry: an hour of data is 360 k windows
    (tens of milliseconds of numpy), and callers repeat geometries (fixed packet sizes, equal trial lengths)."""
    start, stop = _window_indices_cached(int(rows), int(sr), float(window_length), float(window_shift))
    return start, stop


@functools.lru_cache(maxsize=32)
def _window_indices_cached(rows: int, sr: int, window_length: float, window_shift: float):
    start, stop = _window_indices(rows, sr, window_length, window_shift)
    start.setflags(write=False)
    stop.setflags(write=False)
    return start, stop


def _window_indices(rows: int, sr: int, window_length: float, window_shift: float) -> Tuple[np.ndarray, np.ndarray]:
    f32 = np.float32
    wl, ws, srf = f32(window_length), f32(window_shift), f32(sr)
    num = int(np.floor(np.float64((f32(rows) - wl * srf) / (ws * srf)))) + 1
    if num <= 0:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    win = np.arange(num, dtype=np.int64).astype(f32)
    c_round = lambda v: np.copysign(np.floor(np.abs(v) + 0.5), v)  # libm round(): half away from zero
    start = c_round(((win * ws) * srf).astype(np.float64)).astype(np.int64)
    stop = c_round((start.astype(f32) + wl * srf).astype(np.float64)).astype(np.int64)
    return start, stop


def windows_are_regular(start: np.ndarray, stop: np.ndarray, L: int, hop: int, rows: int) -> bool:
    """True when the float32 index arithmetic lands on the integer grid start = w*hop, stop = start + L, the count is
    (rows - L)//hop + 1 and the last window ends inside the data — the case the fused kernels implement."""
    n = len(start)
    if n != ((rows - L) // hop + 1 if rows >= L else 0):
        return False
    if n == 0:
        return True
    w = np.arange(n, dtype=np.int64)
    return bool(np.array_equal(start, w * hop) and np.array_equal(stop, start + L) and stop[-1] <= rows)


@functools.lru_cache(maxsize=32)
def window_plan(rows: int, sr: int, window_length: float, window_shift: float, L: int, hop: int):
    """(start, stop, regular) for a geometry, cached: `window_indices` plus `windows_are_regular`."""
    start, stop = window_indices(rows, sr, window_length, window_shift)
    return start, stop, windows_are_regular(start, stop, L, hop, rows)
