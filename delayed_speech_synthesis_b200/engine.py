"""`HgaEngine`: one libhga_b200 handle = one extractor's device state, plus the handle-less building blocks.

Host arrays are numpy float64; device arrays are torch CUDA float64 tensors (torch is plumbing: device memory and
streams).  Every method ends in a call through the C ABI (include/hga_b200.h); nothing here computes features on the
CPU.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _lib, filters
from ._lib import Config, HgaError, check


def _torch():
    import torch

    return torch


def _stream_ptr(device: int) -> int:
    return int(_torch().cuda.current_stream(device).cuda_stream)


def _vec_i64(a: Sequence[int]):
    arr = np.ascontiguousarray(a, dtype=np.int64)
    return arr, arr.ctypes.data_as(C.POINTER(C.c_int64))


def _host_f64(a, what: str) -> np.ndarray:
    a = np.asarray(a)
    if a.dtype != np.float64:
        raise ValueError(f"{what}: Buffer dtype mismatch, expected 'double' but got '{a.dtype}'")
    if a.ndim != 2:
        raise ValueError(f"{what}: Buffer has wrong number of dimensions (expected 2, got {a.ndim})")
    if a.strides[1] != 8 or a.strides[0] < a.shape[1] * 8 or a.strides[0] % 8:
        a = np.ascontiguousarray(a)
    return a


def _musigma_host(mu, sigma, n_channels: int) -> Tuple[Optional[np.ndarray], Optional[np.ndarray]]:
    if mu is None and sigma is None:
        return None, None
    if mu is None or sigma is None:
        raise ValueError("mu and sigma must be given together")
    mu = np.ascontiguousarray(np.broadcast_to(np.asarray(mu, dtype=np.float64).reshape(-1), (n_channels,)))
    sigma = np.ascontiguousarray(np.broadcast_to(np.asarray(sigma, dtype=np.float64).reshape(-1), (n_channels,)))
    return mu, sigma


class HgaEngine:
    """Device-side state of one `HighGammaExtractor` (reference: local/units.py:97-161)."""

    def __init__(self, fs: int, n_channels: int, window_length: float = 0.05, window_shift: float = 0.01,
                 l_freq: float = 70, h_freq: float = 170, precision: str = "f64", device: Optional[int] = None,
                 warmup: int = 0, warmup_tol: float = 0.0, max_push_rows: int = 0,
                 sos_bp: Optional[np.ndarray] = None, sos_notch: Optional[np.ndarray] = None):
        self._h = None
        lib = _lib.load()
        _lib.require_device()
        if device is None:
            device = _torch().cuda.current_device()
        self.device = int(device)
        self.fs, self.n_channels = int(fs), int(n_channels)
        self.window_length, self.window_shift = window_length, window_shift
        self.frame_len, self.hop = filters.window_geometry(fs, window_length, window_shift)
        if sos_bp is None:
            sos_bp, sos_notch = filters.design_hga_filters(fs, l_freq, h_freq)
        self.sos_bp = np.ascontiguousarray(sos_bp, dtype=np.float64)
        self.sos_notch = np.zeros((0, 6)) if sos_notch is None else np.ascontiguousarray(sos_notch, dtype=np.float64)
        self.zi_bp = filters.initial_state(self.sos_bp)
        self.zi_notch = filters.initial_state(self.sos_notch) if len(self.sos_notch) else np.zeros((0, 2))
        cfg = Config()
        cfg.abi_version, cfg.device, cfg.fs, cfg.n_channels = _lib.HGA_ABI_VERSION, self.device, self.fs, self.n_channels
        cfg.frame_len, cfg.hop, cfg.eps = self.frame_len, self.hop, 0.01
        cfg.n_sections_bp, cfg.n_sections_notch = len(self.sos_bp), len(self.sos_notch)
        for dst, src in ((cfg.sos_bp, self.sos_bp), (cfg.sos_notch, self.sos_notch), (cfg.zi_bp, self.zi_bp),
                         (cfg.zi_notch, self.zi_notch)):
            flat = np.asarray(src, dtype=np.float64).ravel()
            if len(flat) > len(dst):
                raise ValueError("too many filter sections")
            for i, v in enumerate(flat):
                dst[i] = float(v)
        cfg.precision = _lib.PRECISIONS[precision]
        cfg.warmup, cfg.warmup_tol, cfg.max_push_rows = int(warmup), float(warmup_tol), int(max_push_rows)
        self.precision = precision
        h = C.c_void_p()
        check(lib.hga_create(C.byref(cfg), C.byref(h)))
        self._h, self._lib = h, lib
        self.warmup = int(lib.hga_warmup(h))

    # -- life cycle ------------------------------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.hga_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self) -> None:
        check(self._lib.hga_reset(self._h, _stream_ptr(self.device)))

    def num_frames(self, rows: int) -> int:
        return int(self._lib.hga_num_frames(self._h, int(rows)))

    def stream_frames(self, n: int) -> int:
        return int(self._lib.hga_stream_frames(self._h, int(n)))

    def last_plan(self) -> dict:
        """How the last `extract_offline*` call was cut into time-chunks: rows per chunk, chunks, chunks that started from a
        carried filter state, rows the carry reaches."""
        rows, n, nc, k = C.c_int64(0), C.c_int32(0), C.c_int32(0), C.c_int32(0)
        check(self._lib.hga_last_plan(self._h, C.byref(rows), C.byref(n), C.byref(nc), C.byref(k)))
        return {"chunk_rows": rows.value, "chunks": n.value, "carried": nc.value, "carry_rows": k.value}

    def set_timing(self, on: bool) -> None:
        check(self._lib.hga_set_timing(self._h, int(on)))

    def last_kernel_ms(self) -> float:
        return float(self._lib.hga_last_kernel_ms(self._h))

    def last_kernel_split_ms(self):
        """(dominant kernel alone, carry + fix-up kernels behind it) of the last timed device extraction, in ms."""
        a, b = C.c_double(0), C.c_double(0)
        check(self._lib.hga_last_kernel_split_ms(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    # -- offline ---------------------------------------------------------------------------------------------------
    def _segments(self, rows: int, segments):
        if segments is None:
            segments = [(0, rows)]
        start = [int(a) for a, _ in segments]
        length = [int(b) - int(a) for a, b in segments]
        frames = [self.num_frames(n) for n in length]
        out_row = np.concatenate([[0], np.cumsum(frames)])[:-1] if frames else []
        return start, length, frames, [int(v) for v in out_row]

    def extract_offline(self, x, segments=None, mu=None, sigma=None, out=None):
        """Features of independent segments [(start, stop), ...] of x, each from a fresh extractor, concatenated
        (prepare_corpus.py:42-52) with the optional offline z-score (prepare_corpus.py:226-227).

        x: torch CUDA float64 (T, C) -> returns a torch CUDA tensor (enqueued on the current stream);
        x: numpy float64 (T, C)      -> returns numpy (pipelined host<->device copies inside the call)."""
        torch = _torch()
        if isinstance(x, torch.Tensor):
            if not x.is_cuda or x.dtype != torch.float64 or x.dim() != 2 or x.stride(1) != 1:
                raise ValueError("device input must be a CUDA float64 (T, C) tensor with unit channel stride")
            if x.shape[1] != self.n_channels:   # the reference's sosfilt raises on the zi shape mismatch
                raise ValueError(f"expected {self.n_channels} channels, got {x.shape[1]}")
            start, length, frames, out_row = self._segments(x.shape[0], segments)
            total = int(sum(frames))
            if out is None:
                out = torch.empty((total, self.n_channels), dtype=torch.float64, device=x.device)
            mu_t = sg_t = None
            if mu is not None:
                mu_t = torch.as_tensor(mu, dtype=torch.float64, device=x.device).reshape(-1).contiguous()
                sg_t = torch.as_tensor(sigma, dtype=torch.float64, device=x.device).reshape(-1).contiguous()
            (_, ps), (_, pl), (_, po) = a1, a2, a3 = _vec_i64(start), _vec_i64(length), _vec_i64(out_row)
            check(self._lib.hga_extract_offline(self._h, x.data_ptr(), x.stride(0), len(start), ps, pl, po,
                                                mu_t.data_ptr() if mu_t is not None else None,
                                                sg_t.data_ptr() if sg_t is not None else None,
                                                out.data_ptr(), out.stride(0), _stream_ptr(self.device)))
            return out
        x = _host_f64(x, "data")
        if x.shape[1] != self.n_channels:
            raise ValueError(f"expected {self.n_channels} channels, got {x.shape[1]}")
        start, length, frames, out_row = self._segments(x.shape[0], segments)
        total = int(sum(frames))
        if out is None:
            out = np.empty((total, self.n_channels), dtype=np.float64)
        mu_h, sg_h = _musigma_host(mu, sigma, self.n_channels)
        (_, ps), (_, pl), (_, po) = a1, a2, a3 = _vec_i64(start), _vec_i64(length), _vec_i64(out_row)
        check(self._lib.hga_extract_offline_host(self._h, x.ctypes.data, x.strides[0] // 8, len(start), ps, pl, po,
                                                 mu_h.ctypes.data if mu_h is not None else None,
                                                 sg_h.ctypes.data if sg_h is not None else None,
                                                 out.ctypes.data, out.strides[0] // 8))
        return out

    def extract_offline_stats(self, x, stats, segments=None, mu=None, sigma=None, out=None):
        """`extract_offline` of a CUDA float64 (T, C) tensor that also merges the (n, mean, M2) of every frame it writes into
        `stats` (an `ops.ChannelStats`): the day-baseline flow of prepare_corpus.py:190-199 in one pass.  With precision
        'f32' the extraction kernels leave the statistics behind and the features are not read again (`self.last_fused`);
        other handles extract, then accumulate."""
        torch = _torch()
        if not isinstance(x, torch.Tensor) or not x.is_cuda or x.dtype != torch.float64 or x.dim() != 2 or x.stride(1) != 1:
            raise ValueError("device input must be a CUDA float64 (T, C) tensor with unit channel stride")
        if x.shape[1] != self.n_channels or stats.n_channels != self.n_channels or stats.device != self.device:
            raise ValueError(f"expected {self.n_channels} channels on device {self.device}")
        start, length, frames, out_row = self._segments(x.shape[0], segments)
        total = int(sum(frames))
        if out is None:
            out = torch.empty((total, self.n_channels), dtype=torch.float64, device=x.device)
        mu_t = sg_t = None
        if mu is not None:
            mu_t = torch.as_tensor(mu, dtype=torch.float64, device=x.device).reshape(-1).contiguous()
            sg_t = torch.as_tensor(sigma, dtype=torch.float64, device=x.device).reshape(-1).contiguous()
        (_, ps), (_, pl), (_, po) = a1, a2, a3 = _vec_i64(start), _vec_i64(length), _vec_i64(out_row)
        fused = C.c_int32(0)
        check(self._lib.hga_extract_offline_stats(self._h, x.data_ptr(), x.stride(0), len(start), ps, pl, po,
                                                  mu_t.data_ptr() if mu_t is not None else None,
                                                  sg_t.data_ptr() if sg_t is not None else None,
                                                  out.data_ptr(), out.stride(0), stats.acc.data_ptr(), _stream_ptr(self.device), C.byref(fused)))
        self.last_fused = bool(fused.value)
        return out

    def extract_offline_mapped(self, x_raw, chanmap, segments=None, mu=None, sigma=None, out=None):
        """`extract_offline(chanmap(x_raw))` for a CUDA float64 (T, chanmap.spec.c_in) tensor of rows BEFORE the channel map
        (HighGammaExtractor's pre_transform chain, local/units.py:147-152): with precision 'f32' at hop 10 the map runs inside
        the extraction kernel's load stage and the mapped recording is never written.  Sets `self.last_fused`."""
        torch = _torch()
        if not isinstance(x_raw, torch.Tensor) or not x_raw.is_cuda or x_raw.dtype != torch.float64 or x_raw.dim() != 2 or x_raw.stride(1) != 1:
            raise ValueError("raw input must be a CUDA float64 (T, C_in) tensor with unit column stride")
        if x_raw.shape[1] != chanmap.spec.c_in:
            raise ValueError(f"expected {chanmap.spec.c_in} raw columns, got {x_raw.shape[1]}")
        if chanmap.spec.c_out != self.n_channels or chanmap.device != self.device:
            raise ValueError("channel map output width / device do not match the engine")
        start, length, frames, out_row = self._segments(x_raw.shape[0], segments)
        total = int(sum(frames))
        if out is None:
            out = torch.empty((total, self.n_channels), dtype=torch.float64, device=x_raw.device)
        mu_t = sg_t = None
        if mu is not None:
            mu_t = torch.as_tensor(mu, dtype=torch.float64, device=x_raw.device).reshape(-1).contiguous()
            sg_t = torch.as_tensor(sigma, dtype=torch.float64, device=x_raw.device).reshape(-1).contiguous()
        (_, ps), (_, pl), (_, po) = a1, a2, a3 = _vec_i64(start), _vec_i64(length), _vec_i64(out_row)
        fused = C.c_int32(0)
        check(self._lib.hga_extract_offline_mapped(self._h, chanmap._h, x_raw.data_ptr(), x_raw.stride(0), len(start), ps, pl, po,
                                                   mu_t.data_ptr() if mu_t is not None else None,
                                                   sg_t.data_ptr() if sg_t is not None else None,
                                                   out.data_ptr(), out.stride(0), _stream_ptr(self.device), C.byref(fused)))
        self.last_fused = bool(fused.value)
        return out

    def extract_offline_raw(self, raw, chanmap, gain=None, segments=None, mu=None, sigma=None, out=None):
        """`extract_offline` of `chanmap(raw * gain)` for a HOST recording of raw amplifier counts (numpy int16 / int32 /
        float32 / float64, (T, chanmap.spec.c_in)): slabs are uploaded at the raw element size and expanded on the device
        (BCI2000MatFile.signals(), local/common.py:456-459, then the folded pre-transform chain).  chanmap: a
        `transforms.ChannelMap` on this engine's device whose output width is the engine's channel count."""
        from .transforms import RAW_NUMPY

        raw = np.asarray(raw)
        if raw.ndim != 2 or raw.shape[1] != chanmap.spec.c_in:
            raise ValueError(f"expected (T, {chanmap.spec.c_in}) raw data, got {raw.shape}")
        if raw.dtype not in RAW_NUMPY:
            raise TypeError(f"raw element type {raw.dtype} is not one of float64, float32, int32, int16")
        if chanmap.spec.c_out != self.n_channels or chanmap.device != self.device:
            raise ValueError("channel map output width / device do not match the engine")
        if raw.strides[1] != raw.itemsize or raw.strides[0] % raw.itemsize or raw.strides[0] < raw.shape[1] * raw.itemsize:
            raw = np.ascontiguousarray(raw)
        g = None
        if gain is not None:
            g = np.ascontiguousarray(np.broadcast_to(np.asarray(gain, dtype=np.float64).reshape(-1), (raw.shape[1],)))
        start, length, frames, out_row = self._segments(raw.shape[0], segments)
        total = int(sum(frames))
        if out is None:
            out = np.empty((total, self.n_channels), dtype=np.float64)
        mu_h, sg_h = _musigma_host(mu, sigma, self.n_channels)
        (_, ps), (_, pl), (_, po) = a1, a2, a3 = _vec_i64(start), _vec_i64(length), _vec_i64(out_row)
        check(self._lib.hga_extract_offline_raw_host(self._h, chanmap._h, raw.ctypes.data, RAW_NUMPY[raw.dtype],
                                                     g.ctypes.data if g is not None else None, raw.strides[0] // raw.itemsize,
                                                     len(start), ps, pl, po,
                                                     mu_h.ctypes.data if mu_h is not None else None,
                                                     sg_h.ctypes.data if sg_h is not None else None,
                                                     out.ctypes.data, out.strides[0] // 8))
        return out

    # -- streaming -------------------------------------------------------------------------------------------------
    def push(self, x, mu=None, sigma=None):
        """`extract_features` of the long-lived extractor on the next n rows (local/units.py:151-160)."""
        torch = _torch()
        nf = C.c_int64(0)
        if isinstance(x, torch.Tensor):
            if not x.is_cuda or x.dtype != torch.float64 or x.dim() != 2 or x.stride(1) != 1:
                raise ValueError("device input must be a CUDA float64 (n, C) tensor with unit channel stride")
            if x.shape[1] != self.n_channels:
                raise ValueError(f"expected {self.n_channels} channels, got {x.shape[1]}")
            frames = self.stream_frames(x.shape[0])
            out = torch.empty((max(frames, 1), self.n_channels), dtype=torch.float64, device=x.device)[:frames]
            mu_t = sg_t = None
            if mu is not None:
                mu_t = torch.as_tensor(mu, dtype=torch.float64, device=x.device).reshape(-1).contiguous()
                sg_t = torch.as_tensor(sigma, dtype=torch.float64, device=x.device).reshape(-1).contiguous()
            check(self._lib.hga_stream_push(self._h, x.data_ptr(), x.shape[0], x.stride(0),
                                            mu_t.data_ptr() if mu_t is not None else None,
                                            sg_t.data_ptr() if sg_t is not None else None,
                                            out.data_ptr(), max(out.stride(0), self.n_channels), C.byref(nf),
                                            _stream_ptr(self.device)))
            return out
        x = _host_f64(x, "data")
        if x.shape[1] != self.n_channels:
            raise ValueError(f"expected {self.n_channels} channels, got {x.shape[1]}")
        frames = self.stream_frames(x.shape[0])
        out = np.empty((max(frames, 1), self.n_channels), dtype=np.float64)[:frames]
        mu_h, sg_h = self._musigma_cached(mu, sigma)
        check(self._lib.hga_stream_push_host(self._h, x.ctypes.data, x.shape[0], x.strides[0] // 8,
                                             mu_h.ctypes.data if mu_h is not None else None,
                                             sg_h.ctypes.data if sg_h is not None else None,
                                             out.ctypes.data, self.n_channels, C.byref(nf)))
        assert nf.value == frames
        return out

    def _musigma_cached(self, mu, sigma):
        """`_musigma_host`, skipped when the caller passes the same two objects as last time (a streaming extractor hands
        over its ZScoreNormalization's arrays on every packet; the C side compares the contents anyway)."""
        key = getattr(self, "_ms_key", None)
        if key is not None and key[0] is mu and key[1] is sigma:
            return self._ms_val
        val = _musigma_host(mu, sigma, self.n_channels)
        self._ms_key, self._ms_val = (mu, sigma), val
        return val

    def push_packet(self, packet, mu=None, sigma=None, out_f32=None):
        """`interpret_bytes` + `extract_features` for one wire packet (local/units.py:78-82,145-161): bytes of the
        '=BBB HH' header + float32 channel-major payload -> (F, C) float64 numpy frames.  The kernel reads the payload
        as it lies; no (n, C) float64 copy is made.  out_f32: optional CUDA float32 (>= F, C) tensor that also
        receives the frames (for a consumer on the same GPU)."""
        raw = bytes(packet)
        n_ch, n_s = C.c_int32(0), C.c_int32(0)
        check(self._lib.hga_packet_header(raw, len(raw), C.byref(n_ch), C.byref(n_s)))
        frames = self.stream_frames(n_s.value) if n_s.value > 0 else 0
        out = np.empty((max(frames, 1), self.n_channels), dtype=np.float64)[:frames]
        mu_h, sg_h = self._musigma_cached(mu, sigma)
        f32_ptr = None
        if out_f32 is not None:
            torch = _torch()
            if not (isinstance(out_f32, torch.Tensor) and out_f32.is_cuda and out_f32.dtype == torch.float32 and out_f32.is_contiguous()
                    and out_f32.dim() == 2 and out_f32.shape[1] == self.n_channels and out_f32.shape[0] >= frames):
                raise ValueError("out_f32 must be a contiguous CUDA float32 (>= F, C) tensor")
            f32_ptr = out_f32.data_ptr()
        nf = C.c_int64(0)
        check(self._lib.hga_stream_push_packet(self._h, raw, len(raw), mu_h.ctypes.data if mu_h is not None else None,
                                               sg_h.ctypes.data if sg_h is not None else None, out.ctypes.data, self.n_channels,
                                               f32_ptr, C.byref(nf)))
        assert nf.value == frames
        return out

    def get_state(self):
        """(hg_state (S,2,C), fh_state (S,2,C), remainder (L-hop, C), first_frame) in the reference's shapes."""
        Cn = self.n_channels
        hg = np.empty((len(self.sos_bp), 2, Cn))
        fh = np.empty((len(self.sos_notch), 2, Cn))
        rem = np.empty((self.frame_len - self.hop, Cn))
        ff = C.c_int32(0)
        check(self._lib.hga_get_state(self._h, hg.ctypes.data, fh.ctypes.data if fh.size else None,
                                      rem.ctypes.data if rem.size else None, C.byref(ff)))
        return hg, fh, rem, bool(ff.value)

    def set_state(self, hg=None, fh=None, remainder=None, first_frame=None) -> None:
        def prep(a, shape):
            if a is None:
                return None, None
            a = np.ascontiguousarray(a, dtype=np.float64)
            if a.shape != shape:
                raise ValueError(f"state shape {a.shape} != {shape}")
            return a, a.ctypes.data
        Cn = self.n_channels
        hg, p1 = prep(hg, (len(self.sos_bp), 2, Cn))
        fh, p2 = prep(fh, (len(self.sos_notch), 2, Cn))
        rem, p3 = prep(remainder, (self.frame_len - self.hop, Cn))
        ff = None if first_frame is None else C.c_int32(int(bool(first_frame)))
        check(self._lib.hga_set_state(self._h, p1, p2, p3, C.byref(ff) if ff is not None else None))
