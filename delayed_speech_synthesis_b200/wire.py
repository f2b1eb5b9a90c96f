"""Wire packets at the streaming seam, on the GPU.

The reference's `ZMQConnector` (/root/reference/local/units.py:45-93) receives BCI2000 GenericSignal packets: a 7-byte
header `struct '=BBB HH'` (descriptor, supplement, dtype, n_channels, n_samples) and `n_channels * n_samples` float32,
channel-major; `interpret_bytes` (:78-82) turns one into a `(n_samples, n_channels)` float64 array.  Here the decode is
a device kernel (`hga_packet_decode`), or disappears into the extraction kernel (`HgaEngine.push_packet`,
`HighGammaExtractor.extract_features_packet`).
"""
from __future__ import annotations

import ctypes as C
import struct
from typing import Tuple

import numpy as np

from . import _lib
from ._lib import check

HEADER = struct.Struct('=BBB HH')          # local/units.py:48
TOPIC = struct.Struct('=BBB').pack(4, 1, 2)  # local/units.py:62: the subscription prefix of signal packets


def packet_shape(data: bytes) -> Tuple[int, int]:
    """(n_channels, n_samples) of a packet; raises HgaError when the payload length does not match the header."""
    n_ch, n_s = C.c_int32(0), C.c_int32(0)
    check(_lib.load().hga_packet_header(data, len(data), C.byref(n_ch), C.byref(n_s)))
    return n_ch.value, n_s.value


def pack_packet(payload: np.ndarray, descriptor: int = 4, supplement: int = 1, dtype_code: int = 2) -> bytes:
    """(n_channels, n_samples) array -> packet bytes, the layout development_amplifier.py:14-25 sends (test helper and
    synthetic amplifier for the latency bench)."""
    payload = np.ascontiguousarray(payload, dtype=np.float32)
    return HEADER.pack(descriptor, supplement, dtype_code, payload.shape[0], payload.shape[1]) + payload.tobytes()


def interpret_bytes(data: bytes, device=None, as_numpy: bool = False):
    """`ZMQConnector.interpret_bytes` (local/units.py:78-82): packet -> (n_samples, n_channels) float64, C order.
    Returns a torch CUDA tensor (the consumer is on the GPU); `as_numpy=True` copies it back like the reference returns."""
    import torch

    lib = _lib.load()
    _lib.require_device()
    n_ch, n_s = packet_shape(data)
    dev = torch.cuda.current_device() if device is None else int(device)
    x = torch.empty((n_s, n_ch), dtype=torch.float64, device=f"cuda:{dev}")
    if n_ch and n_s:
        with torch.cuda.device(dev):
            check(lib.hga_packet_decode(dev, data, len(data), x.data_ptr(), n_ch, int(torch.cuda.current_stream(dev).cuda_stream)))
            torch.cuda.current_stream(dev).synchronize()   # the packet bytes are pageable: the copy must finish before they go away
    return x.cpu().numpy() if as_numpy else x
