"""GPU-backed stand-in for the reference's compiled extension module `hga_optimized`
(/root/reference/extensions/hga/hga_optimized.pyx): same two symbols, same signatures, dtypes and shapes.

    from delayed_speech_synthesis_b200.hga_optimized import compute_log_power_features, WarmStartFrameBuffer
"""
from __future__ import annotations

import numpy as np

from . import filters, ops
from .engine import _torch


def compute_log_power_features(data, sr: int, window_length: float, window_shift: float):
    """hga_optimized.pyx:27-47 — `log(mean(x^2 over window) + 0.01)` per window (rows) and channel (columns).

    `data`: 2-D float64, numpy (any strides; returns a fresh numpy (F, C) array as the reference does) or a torch CUDA
    tensor (returns a CUDA tensor).  Wrong dtype / ndim raise ValueError like Cython's buffer acquisition.
    """
    return ops.log_power(data, int(sr), window_length, window_shift)


class WarmStartFrameBuffer:
    """hga_optimized.pyx:50-131 — keeps the last (frame_length - frame_shift) rows between calls; a first chunk shorter
    than one frame is zero-padded in front.  Pure data movement (no arithmetic): numpy in -> numpy out on the host,
    CUDA tensor in -> CUDA tensor out on the device."""

    def __init__(self, frame_length: float, frame_shift: float, fs: int, nb_channels: int):
        self.frame_length_in_samples, shift = filters.window_geometry(fs, frame_length, frame_shift)   # pyx:76-77
        self.overlap = self.frame_length_in_samples - shift                                               # pyx:78
        self.first_frame = True                                                                           # pyx:79
        self.nb_channels = nb_channels
        self.remainder_data = np.zeros((self.overlap, nb_channels), dtype=np.float64)                     # pyx:82-88

    def reset(self) -> None:                                                                              # pyx:90-94
        self.first_frame = True
        self.remainder_data = np.zeros((self.overlap, self.nb_channels), dtype=np.float64)

    def _like(self, data, rows: int, zero: bool):
        torch = _torch()
        if isinstance(data, torch.Tensor):
            f = torch.zeros if zero else torch.empty
            return f((rows, data.shape[1]), dtype=torch.float64, device=data.device)
        return (np.zeros if zero else np.empty)((rows, data.shape[1]), dtype=np.float64)

    def insert(self, data):                                                                               # pyx:96-131
        torch = _torch()
        if not isinstance(data, torch.Tensor):
            data = np.asarray(data)
            if data.dtype != np.float64:
                raise ValueError(f"Buffer dtype mismatch, expected 'double' but got '{data.dtype}'")
        if data.ndim != 2:
            raise ValueError(f"Buffer has wrong number of dimensions (expected 2, got {data.ndim})")
        n, L, ov = data.shape[0], self.frame_length_in_samples, self.overlap
        if self.first_frame and n >= L:                                           # case 1: returns the caller's array
            self.first_frame = False
            self.remainder_data = data[n - ov:, :]
            return data
        if self.first_frame:                                                      # case 2
            out = self._like(data, L, zero=True)
            out[L - n:, :] = data
            self.first_frame = False
        else:                                                                     # case 3
            out = self._like(data, ov + n, zero=False)
            rem = self.remainder_data
            if isinstance(out, torch.Tensor) and not isinstance(rem, torch.Tensor):
                rem = torch.from_numpy(np.ascontiguousarray(rem)).to(out.device)
            elif not isinstance(out, torch.Tensor) and isinstance(rem, torch.Tensor):
                rem = rem.cpu().numpy()
            out[:ov, :] = rem
            out[ov:, :] = data
        self.remainder_data = out[out.shape[0] - ov:, :]
        return out
