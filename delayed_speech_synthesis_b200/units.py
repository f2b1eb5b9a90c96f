"""Host-side mirror of the HGA part of the reference's `local/units.py` (lines 93-207): `HighGammaExtractor`,
`HighGammaActivitySettings`, `HighGammaActivityState`, `HighGammaActivity` — same names, constructor arguments,
attributes, message interface and error behaviour, with the arithmetic running in libhga_b200's CUDA kernels.

    from delayed_speech_synthesis_b200.units import HighGammaExtractor, HighGammaActivity, HighGammaActivitySettings
"""
from __future__ import annotations

import logging
from dataclasses import replace
from functools import reduce
from typing import AsyncGenerator, Callable, List, Optional

import numpy as np

from . import filters, transforms
from .common import ZScoreNormalization
from .engine import HgaEngine, _torch
from .ezmsg_compat import TimeSeriesMessage, ez

logger = logging.getLogger("units.py")

Transforms = Optional[List[Callable]]


class _DeviceFrameBuffer:
    """`extractor.framebuffer` (local/units.py:112-113): the warm-start buffer lives in the engine's device state; this
    proxy exposes the reference object's `reset()`."""

    def __init__(self, engine: HgaEngine):
        self._engine = engine

    def reset(self) -> None:                      # hga_optimized.pyx:90-94
        hg, fh, rem, _ = self._engine.get_state()
        self._engine.set_state(remainder=np.zeros_like(rem), first_frame=True)


def _is_zscore(t) -> bool:
    return isinstance(t, ZScoreNormalization) or (
        type(t).__name__ == "ZScoreNormalization" and hasattr(t, "channel_means") and hasattr(t, "channel_stds"))


class HighGammaExtractor:
    """local/units.py:97-161.  `precision`/`device` are additions with defaults; everything else is the reference's.

    precision='f64'  : small pushes (<= 1024 rows) run the reference-order streaming kernel; larger ones the fused,
                       time-chunked float64 kernel (|feature error| <= 1e-9).
    precision='exact': every push keeps the reference's operation order.
    """

    def __init__(self, fs, nb_electrodes, window_length=0.05, window_shift=0.01, l_freq: int = 70, h_freq: int = 170,
                 pre_transforms: Optional[List[Callable]] = None, post_transforms: Optional[List[Callable]] = None,
                 precision: str = "f64", device: Optional[int] = None):
        self.fs = fs
        self.nb_electrodes = nb_electrodes
        self.window_length = window_length
        self.window_shift = window_shift
        self.model_order = 4                                                        # :108 (unused there as well)
        self.step_size = 5                                                          # :109
        self.pre_transform = pre_transforms
        self.post_transform = post_transforms

        # A ZScoreNormalization that is the ONLY post-transform is folded into the feature kernel
        # (decode_online.py:94-95,109).  Chains of column selections / common average referencing / bad-channel
        # correction (prepare_corpus.py:152-169, decode_online.py:66-82) are folded into one channel-map kernel each
        # (transforms.fold); any other chain is applied as the reference does, by calling the user's callables in order.
        self._fused_zscore = None
        self._pre_list = list(pre_transforms) if pre_transforms is not None else None
        self._pre_map = None          # (input width, transforms.ChannelMap or None), built on the first call
        self._post_map = None
        if post_transforms is not None and len(post_transforms) == 1 and _is_zscore(post_transforms[0]):
            self._fused_zscore = post_transforms[0]
        elif post_transforms:
            spec = transforms.fold(post_transforms, nb_electrodes)
            if spec is not None and spec.c_out == nb_electrodes:
                self._post_map = spec
        if self.pre_transform is not None:
            self.pre_transform = self._compose_functions(*pre_transforms)          # :115-116
        if self.post_transform is not None:
            self.post_transform = self._compose_functions(*post_transforms)        # :117-118

        if not ((60 < l_freq < 120) or (120 < h_freq < 180)):                      # :120-121
            logger.warning('l_freq and h_freq seem not to be in the recommended ranges!!')

        self.hg_filter = filters.create_filter_sos(fs, l_freq, h_freq)             # :124-125
        self.fh_filter = filters.create_filter_sos(fs, 122, 118)                   # :126
        self._engine = HgaEngine(fs, nb_electrodes, window_length, window_shift, precision=precision, device=device,
                                 sos_bp=self.hg_filter, sos_notch=self.fh_filter)
        self.framebuffer = _DeviceFrameBuffer(self._engine)                        # :112-113

    # filter states in the reference's (8, 2, C) shape, read from / written to the device (local/units.py:128-132)
    @property
    def hg_state(self) -> np.ndarray:
        return self._engine.get_state()[0]

    @hg_state.setter
    def hg_state(self, value) -> None:
        self._engine.set_state(hg=value)

    @property
    def fh_state(self) -> np.ndarray:
        return self._engine.get_state()[1]

    @fh_state.setter
    def fh_state(self, value) -> None:
        self._engine.set_state(fh=value)

    @staticmethod
    def _compose_functions(*functions):                                            # :134-136
        return reduce(lambda f, g: lambda x: g(f(x)), functions, lambda x: x)

    @staticmethod
    def create_filter(sr, l_freq, h_freq, method='fir', iir_params=None):          # :138-143
        """Dict with the designed filter under "sos", like mne's `create_filter` for method='iir'."""
        if method != 'iir':
            raise NotImplementedError("only method='iir' is on the HGA path")
        order = (iir_params or {}).get('order', 8)
        return {"sos": filters.create_filter_sos(sr, l_freq, h_freq, order), "order": order, "ftype": "butter"}

    def _apply_pre(self, data):
        """pre_transforms (:147-148): one channel-map kernel when the chain folds, else the callables themselves."""
        torch = _torch()
        width = int(data.shape[1]) if getattr(data, "ndim", 0) == 2 else -1
        if width > 0 and (self._pre_map is None or self._pre_map[0] != width):
            spec = transforms.fold(self._pre_list, width)
            ok = spec is not None and spec.c_out == self.nb_electrodes
            self._pre_map = (width, transforms.ChannelMap(spec, self._engine.device) if ok else None)
        cmap = self._pre_map[1] if self._pre_map is not None and width > 0 else None
        if cmap is None:
            return self.pre_transform(data)
        if not isinstance(data, torch.Tensor):
            data = np.ascontiguousarray(data, dtype=np.float64)
            if data.shape[0] * data.shape[1] >= (1 << 16):      # large blocks stay on the device for the extraction
                data = torch.from_numpy(data).cuda(self._engine.device)
                return ("numpy", cmap.apply(data))
        return cmap.apply(data)

    def extract_features_packet(self, packet: bytes):
        """`ZMQConnector.interpret_bytes` (:78-82) followed by `extract_features` (:145-161) for one wire packet.  Without
        pre-transforms the extraction kernel reads the packet's float32 channel-major payload directly; with them the
        packet is decoded on the device and the (n, C_in) tensor takes the usual route."""
        if self.pre_transform is None and self._post_map is None and (self.post_transform is None or self._fused_zscore is not None):
            mu = sigma = None
            if self._fused_zscore is not None:
                mu, sigma = self._fused_zscore.channel_means, self._fused_zscore.channel_stds
            return self._engine.push_packet(packet, mu, sigma)
        from . import wire
        n_ch, _ = wire.packet_shape(packet)
        spec = transforms.fold(self._pre_list, n_ch) if self._pre_list else None
        on_device = spec is not None and spec.c_out == self.nb_electrodes      # else the user's callables want ndarrays
        out = self.extract_features(wire.interpret_bytes(packet, self._engine.device, as_numpy=not on_device))
        return out.cpu().numpy() if isinstance(out, _torch().Tensor) else out

    def extract_features(self, data):                                              # :145-161
        as_numpy = False
        if self.pre_transform is not None:
            data = self._apply_pre(data)
            if isinstance(data, tuple):
                as_numpy, data = True, data[1]
        mu = sigma = None
        if self._fused_zscore is not None:
            mu, sigma = self._fused_zscore.channel_means, self._fused_zscore.channel_stds
        torch = _torch()
        if not isinstance(data, torch.Tensor):
            data = np.asarray(data)
            if data.dtype != np.float64:   # scipy.signal.sosfilt upcasts (float32 wire data becomes float64, :81)
                data = data.astype(np.float64)
        data = self._engine.push(data, mu, sigma)
        if self._post_map is not None:
            if not isinstance(self._post_map, transforms.ChannelMap):
                self._post_map = transforms.ChannelMap(self._post_map, self._engine.device)
            data = self._post_map.apply(data)
        elif self.post_transform is not None and self._fused_zscore is None:
            data = self.post_transform(data)
        if as_numpy and isinstance(data, torch.Tensor):
            data = data.cpu().numpy()
        return data


class HighGammaActivitySettings(ez.Settings):
    """local/units.py:164-175.  Window length and window shift in seconds."""
    fs: int
    nb_electrodes: int
    window_length: float = 0.05
    window_shift: float = 0.01
    l_freq: int = 70
    h_freq: int = 170
    pre_transforms: Transforms = None
    post_transforms: Transforms = None


class HighGammaActivityState(ez.State):
    """local/units.py:178-182."""
    hg_extractor: Optional[HighGammaExtractor] = None


class HighGammaActivity(ez.Unit):
    """local/units.py:185-207: extracts 70-170 Hz high-gamma log-power features, removing the first line-noise
    harmonic; emits one message per input message with `fs = 1 / window_shift`."""
    SETTINGS: HighGammaActivitySettings
    STATE: HighGammaActivityState

    INPUT = ez.InputStream(TimeSeriesMessage)
    OUTPUT = ez.OutputStream(TimeSeriesMessage)

    def initialize(self) -> None:
        # like the reference (:196-201), l_freq / h_freq of the settings are NOT forwarded
        self.STATE.hg_extractor = HighGammaExtractor(
            fs=self.SETTINGS.fs, nb_electrodes=self.SETTINGS.nb_electrodes,
            window_length=self.SETTINGS.window_length, window_shift=self.SETTINGS.window_shift,
            pre_transforms=self.SETTINGS.pre_transforms, post_transforms=self.SETTINGS.post_transforms
        )

    @ez.publisher(OUTPUT)
    @ez.subscriber(INPUT)
    async def process(self, msg: TimeSeriesMessage) -> AsyncGenerator:
        features = self.STATE.hg_extractor.extract_features(msg.data)
        yield self.OUTPUT, replace(msg, data=features, fs=1 / self.SETTINGS.window_shift)
