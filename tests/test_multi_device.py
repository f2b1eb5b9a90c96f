"""One process holding handles on two devices (the C ABI takes a device ordinal per handle): kernels that opt in to more than
48 KB of dynamic shared memory must do so on every device they run on.  Needs two visible GPUs; skipped otherwise."""
from __future__ import annotations

import numpy as np
import pytest

from oracle import hga_oracle as O

pytestmark = pytest.mark.gpu


def _two_devices():
    import torch
    return torch.cuda.is_available() and torch.cuda.device_count() >= 2


def test_handles_on_two_devices_agree_bit_for_bit():
    if not _two_devices():
        pytest.skip("needs two visible GPUs")
    import torch
    from delayed_speech_synthesis_b200 import transforms as TR
    from delayed_speech_synthesis_b200.engine import HgaEngine

    fs, C, T = 1000, 128, 60_000
    x = O.synthetic_ecog(T, C, fs, "ecog", session=9)
    mu, sigma = np.linspace(1.0, 2.0, C), np.linspace(0.5, 1.5, C)
    outs = {}
    for prec in ("f64", "f32"):
        for dev in (0, 1):
            eng = HgaEngine(fs, C, precision=prec, device=dev)
            outs[prec, dev] = eng.extract_offline(x, None, mu, sigma)                       # host buffers
            xd = torch.from_numpy(x).cuda(dev)
            with torch.cuda.device(dev):
                d = eng.extract_offline(xd, None, mu, sigma)                                # device buffers
                torch.cuda.synchronize(dev)
            assert np.array_equal(d.cpu().numpy(), outs[prec, dev])
        assert np.array_equal(outs[prec, 0], outs[prec, 1])
    ref = O.HighGammaExtractor(fs, C, post_transforms=[O.ZScoreNormalization(mu, sigma)]).extract_features(x[:5000])
    assert np.abs(outs["f64", 1][:len(ref)] - ref).max() <= 1e-9
    # a channel map wide enough to need the shared-memory opt-in, on both devices
    wide = np.random.Generator(np.random.Philox(key=[20241104, 77])).standard_normal((500, 1024))
    sel = np.arange(1023, -1, -1)
    for dev in (0, 1):
        m = TR.ChannelMap(TR.fold([TR.ChannelSelection(sel)], 1024), dev)
        got = m.apply(torch.from_numpy(wide).cuda(dev))
        assert np.array_equal(got.cpu().numpy(), wide[:, sel])
    # streaming on the second device
    eng = HgaEngine(fs, C, device=1)
    o = O.HighGammaExtractor(fs, C)
    for t in range(0, 300, 10):
        a, b = eng.push(x[t:t + 10]), o.extract_features(x[t:t + 10])
        assert a.shape == b.shape and np.abs(np.asarray(a) - b).max() <= 1e-12
