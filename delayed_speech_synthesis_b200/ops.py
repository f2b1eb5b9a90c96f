"""Handle-less building blocks over the C ABI: sosfilt, windowed log-power, z-score, Chan-merge statistics.
Inputs may be numpy float64 (copied to the device and back) or torch CUDA float64 tensors (stay on the device)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np

from . import _lib, filters
from ._lib import check
from .engine import _host_f64, _stream_ptr, _torch


def _to_device(a, device=None):
    torch = _torch()
    _lib.require_device()
    if isinstance(a, torch.Tensor):
        if not a.is_cuda:
            a = a.cuda(device)
        return a if a.dtype == torch.float64 else a.double()
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda(device)


def _dev_index(t) -> int:
    return t.device.index if t.device.index is not None else _torch().cuda.current_device()


def sosfilt(sos: np.ndarray, x, zi) -> Tuple[object, object]:
    """`scipy.signal.sosfilt(sos, x, axis=0, zi=zi)` on the GPU, bit-identical arithmetic (local/units.py:151-152).
    x: (T, C); zi: (n_sections, 2, C).  Returns (y, zf) as numpy if x was numpy, else torch CUDA tensors."""
    torch = _torch()
    as_numpy = not isinstance(x, torch.Tensor)
    xd = _to_device(_host_f64(x, "x") if as_numpy else x)
    if xd.dim() != 2 or xd.stride(1) != 1:
        xd = xd.contiguous()
    sos = np.ascontiguousarray(sos, dtype=np.float64)
    T, Cn = xd.shape
    state = _to_device(zi, xd.device).clone().contiguous()
    if tuple(state.shape) != (sos.shape[0], 2, Cn):
        raise ValueError(f"zi shape {tuple(state.shape)} != {(sos.shape[0], 2, Cn)}")
    y = torch.empty((T, Cn), dtype=torch.float64, device=xd.device)
    dev = _dev_index(xd)
    with torch.cuda.device(dev):
        check(_lib.load().hga_sosfilt(dev, sos.ctypes.data, sos.shape[0], xd.data_ptr(), T, Cn, max(xd.stride(0), Cn),
                                      state.data_ptr(), y.data_ptr(), Cn, _stream_ptr(dev)))
    if as_numpy:
        return y.cpu().numpy(), state.cpu().numpy()
    return y, state


def log_power(data, sr: int, window_length: float, window_shift: float, mu=None, sigma=None, zero_rows: int = 0,
              eps: float = 0.01):
    """compute_log_power_features (hga_optimized.pyx:27-47) with an optional fused z-score.  Window indices follow the
    reference's float32 arithmetic; irregular index sets are passed to the kernel explicitly."""
    torch = _torch()
    as_numpy = not isinstance(data, torch.Tensor)
    d = _to_device(_host_f64(data, "data") if as_numpy else data)
    if d.dim() != 2:
        raise ValueError(f"Buffer has wrong number of dimensions (expected 2, got {d.dim()})")
    if d.stride(1) != 1:
        d = d.contiguous()
    rows, Cn = d.shape
    L, hop = filters.window_geometry(sr, window_length, window_shift)
    start, stop, regular = filters.window_plan(int(rows + zero_rows), int(sr), float(window_length), float(window_shift), int(L), int(hop))
    F = len(start)
    dev = _dev_index(d)
    out = torch.empty((F, Cn), dtype=torch.float64, device=d.device)
    if F == 0:
        return out.cpu().numpy() if as_numpy else out
    sd = ed = None
    if not regular:
        if F and (start.min() < 0 or stop.max() > rows + zero_rows):
            raise ValueError("window indices fall outside the data (the reference would read out of bounds)")
        sd, ed = torch.from_numpy(start.copy()).to(d.device), torch.from_numpy(stop.copy()).to(d.device)
    mu_t = sg_t = None
    if mu is not None:
        mu_t = torch.as_tensor(mu, dtype=torch.float64, device=d.device).reshape(-1).expand(Cn).contiguous()
        sg_t = torch.as_tensor(sigma, dtype=torch.float64, device=d.device).reshape(-1).expand(Cn).contiguous()
    with torch.cuda.device(dev):
        check(_lib.load().hga_log_power(dev, d.data_ptr() if rows else None, rows, zero_rows, Cn, max(d.stride(0), Cn), F,
                                        L, hop, sd.data_ptr() if sd is not None else None,
                                        ed.data_ptr() if ed is not None else None, eps,
                                        mu_t.data_ptr() if mu_t is not None else None,
                                        sg_t.data_ptr() if sg_t is not None else None, out.data_ptr(), Cn,
                                        _stream_ptr(dev)))
    return out.cpu().numpy() if as_numpy else out


def zscore(features, mu, sigma):
    """(features - mu) / sigma (local/common.py:375-376; prepare_corpus.py:226-227)."""
    torch = _torch()
    as_numpy = not isinstance(features, torch.Tensor)
    f = _to_device(features)
    if f.dim() != 2 or f.stride(1) != 1:
        f = f.reshape(-1, f.shape[-1]).contiguous()
    F, Cn = f.shape
    dev = _dev_index(f)
    mu_t = torch.as_tensor(mu, dtype=torch.float64, device=f.device).reshape(-1).expand(Cn).contiguous()
    sg_t = torch.as_tensor(sigma, dtype=torch.float64, device=f.device).reshape(-1).expand(Cn).contiguous()
    out = torch.empty((F, Cn), dtype=torch.float64, device=f.device)
    if F == 0:
        return out.cpu().numpy() if as_numpy else out
    with torch.cuda.device(dev):
        check(_lib.load().hga_zscore(dev, f.data_ptr(), F, Cn, max(f.stride(0), Cn), mu_t.data_ptr(), sg_t.data_ptr(),
                                     out.data_ptr(), Cn, _stream_ptr(dev)))
    return out.cpu().numpy() if as_numpy else out


class ChannelStats:
    """Per-channel (n, mean, M2) accumulator on the device: the day-baseline statistics of prepare_corpus.py:190-199 /
    baseline_offline.py:45-60 as a mergeable reduction (Chan et al.)."""

    def __init__(self, n_channels: int, device: Optional[int] = None):
        torch = _torch()
        _lib.require_device()
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.n_channels = int(n_channels)
        self.acc = torch.zeros((3, self.n_channels), dtype=torch.float64, device=f"cuda:{self.device}")

    def update(self, features) -> "ChannelStats":
        torch = _torch()
        f = _to_device(features, self.device)
        if f.dim() != 2 or f.shape[1] != self.n_channels:
            raise ValueError(f"expected (F, {self.n_channels}) features, got {tuple(f.shape)}")
        if f.stride(1) != 1:
            f = f.contiguous()
        if f.shape[0] == 0:
            return self
        with torch.cuda.device(self.device):
            check(_lib.load().hga_stats_accumulate(self.device, f.data_ptr(), f.shape[0], self.n_channels,
                                                   max(f.stride(0), self.n_channels), self.acc.data_ptr(),
                                                   _stream_ptr(self.device)))
        return self

    def merge(self, others) -> "ChannelStats":
        """Merge accumulators given as a (k, 3, C) tensor/array (or another ChannelStats), in index order."""
        torch = _torch()
        if isinstance(others, ChannelStats):
            others = others.acc
        b = _to_device(others, self.device).reshape(-1, 3, self.n_channels).contiguous()
        with torch.cuda.device(self.device):
            check(_lib.load().hga_stats_merge(self.device, self.acc.data_ptr(), b.data_ptr(), b.shape[0], self.n_channels,
                                              _stream_ptr(self.device)))
        return self

    def finalize(self, as_numpy: bool = True):
        """(mean, population std) — np.mean / np.std(ddof=0) over all frames seen."""
        torch = _torch()
        mu = torch.empty(self.n_channels, dtype=torch.float64, device=self.acc.device)
        sg = torch.empty_like(mu)
        with torch.cuda.device(self.device):
            check(_lib.load().hga_stats_finalize(self.device, self.acc.data_ptr(), self.n_channels, mu.data_ptr(),
                                                 sg.data_ptr(), _stream_ptr(self.device)))
        return (mu.cpu().numpy(), sg.cpu().numpy()) if as_numpy else (mu, sg)
