"""Host-side mirror of the `local/common.py` classes on and next to the HGA path: `ZScoreNormalization` here, the
channel-axis transforms (`ChannelSelection`, `CommonAverageReferencing`, `BadChannelCorrection`) in transforms.py."""
from __future__ import annotations

import numpy as np

from .transforms import BadChannelCorrection, ChannelSelection, CommonAverageReferencing  # noqa: F401  (local/common.py:16-58,220-365)


class ZScoreNormalization:
    """Normalizes each channel with predefined means and standard deviations (/root/reference/local/common.py:367-376).

    As a `post_transforms` entry of `HighGammaExtractor` it is recognised and fused into the feature kernel
    (decode_online.py:94-95,109).  Called directly it runs the z-score kernel on the GPU; there is no CPU path."""

    def __init__(self, channel_means: np.ndarray, channel_stds: np.ndarray):
        self.channel_means = channel_means
        self.channel_stds = channel_stds

    def __call__(self, data):
        from . import ops

        arr = data if hasattr(data, "is_cuda") else np.asarray(data, dtype=np.float64)
        shape = arr.shape
        out = ops.zscore(arr.reshape(-1, shape[-1]) if arr.ndim != 2 else arr, self.channel_means, self.channel_stds)
        return out.reshape(shape)
