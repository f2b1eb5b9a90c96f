"""Wire packets at the streaming seam: '=BBB HH' header + float32 channel-major payload.

CPU part: the oracle's restatements of ZMQConnector.interpret_bytes (/root/reference/local/units.py:78-82) and
BCI2000Package.pack (development_amplifier.py:14-25) against vectors produced by the reference's own code
(tests/golden/make_golden.py packets), and the host-side header validation of the C ABI (no device needed).
GPU part: the device decode and the packet-fed extraction kernel through the C ABI against the oracle.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import pytest

from oracle import hga_oracle as O
from delayed_speech_synthesis_b200 import HgaError, wire

G = np.load(Path(__file__).parent / "golden" / "packets.npz")
NAMES = ["a", "b", "c", "d"]


def _pkt(name):
    return G[f"packet_{name}"].tobytes()


# ------------------------------------------------------------------------------------------------------ oracle pins
@pytest.mark.parametrize("name", NAMES)
def test_oracle_interpret_bytes_matches_reference(name):
    got = O.interpret_bytes(_pkt(name))
    want = G[f"decoded_{name}"]
    assert got.dtype == np.float64 and got.flags.c_contiguous and got.shape == want.shape
    assert np.array_equal(got, want)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_pack_matches_reference_packer(name):
    payload = G[f"decoded_{name}"].T.astype(np.float32)
    assert O.pack_packet(payload) == _pkt(name)
    assert wire.pack_packet(payload) == _pkt(name)


# ------------------------------------------------------------------------------------------------------ host logic
def test_header_parse_and_validation():
    assert wire.packet_shape(_pkt("a")) == (128, 10)
    assert wire.packet_shape(_pkt("d")) == (1, 1)
    assert wire.HEADER.size == 7 and wire.TOPIC == bytes([4, 1, 2])
    with pytest.raises(HgaError):
        wire.packet_shape(_pkt("a")[:-4])        # payload shorter than the header says
    with pytest.raises(HgaError):
        wire.packet_shape(_pkt("a") + b"\0")      # longer
    with pytest.raises(HgaError):
        wire.packet_shape(b"\x04\x01\x02")        # shorter than the header
    assert wire.packet_shape(wire.HEADER.pack(4, 1, 2, 0, 0)) == (0, 0)


# ------------------------------------------------------------------------------------------------------ GPU, C ABI
@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_device_decode_is_bit_exact(name):
    x = wire.interpret_bytes(_pkt(name))
    assert x.is_cuda and x.shape == G[f"decoded_{name}"].shape
    assert np.array_equal(x.cpu().numpy(), G[f"decoded_{name}"])
    assert np.array_equal(wire.interpret_bytes(_pkt(name), as_numpy=True), G[f"decoded_{name}"])


@pytest.mark.gpu
def test_device_decode_ragged_shapes():
    rng = np.random.Generator(np.random.Philox(key=[20241104, 12]))
    for c, n in [(33, 65), (1024, 20), (7, 1000), (0, 0)]:
        pkt = O.pack_packet((rng.standard_normal((c, n)) * 30).astype(np.float32))
        assert np.array_equal(wire.interpret_bytes(pkt, as_numpy=True), O.interpret_bytes(pkt))


@pytest.mark.gpu
def test_packet_stream_matches_oracle_and_array_path():
    """decode_online.py wiring: ZMQ packets of 10 samples -> extractor with the day's z-score; packet-fed kernel ==
    array-fed kernel bit for bit, both within the exact mode's tolerance of the oracle."""
    import torch
    from delayed_speech_synthesis_b200.common import ZScoreNormalization
    from delayed_speech_synthesis_b200.units import HighGammaExtractor

    fs, C, n = 1000, 128, 10
    x32 = O.synthetic_ecog(400, C, fs, "ecog", session=3).astype(np.float32)   # the wire carries float32
    mu, sigma = np.linspace(2.0, 3.0, C), np.linspace(0.5, 1.5, C)
    ex_p = HighGammaExtractor(fs, C, post_transforms=[ZScoreNormalization(mu, sigma)])
    ex_a = HighGammaExtractor(fs, C, post_transforms=[ZScoreNormalization(mu, sigma)])
    ref = O.HighGammaExtractor(fs, C, post_transforms=[O.ZScoreNormalization(mu, sigma)])
    f32 = torch.zeros((4, C), dtype=torch.float32, device="cuda")
    for t in range(0, 400, n):
        pkt = O.pack_packet(x32[t:t + n].T)
        a = ex_p.extract_features_packet(pkt)
        b = ex_a.extract_features(O.interpret_bytes(pkt))
        want = ref.extract_features(O.interpret_bytes(pkt))
        assert a.shape == b.shape == want.shape == (1, C)
        assert np.array_equal(a, b)
        assert np.abs(a - want).max() <= 1e-12
    # float32 emission for a consumer on the same GPU
    eng = ex_p._engine
    pkt = O.pack_packet(x32[:n].T)
    frames = eng.push_packet(pkt, mu, sigma, out_f32=f32)
    assert np.array_equal(f32[:1].cpu().numpy(), frames.astype(np.float32))
    with pytest.raises(HgaError):
        eng.push_packet(O.pack_packet(x32[:n, :64].T))          # wrong channel count


@pytest.mark.gpu
def test_packet_with_folded_pre_transforms():
    from delayed_speech_synthesis_b200 import transforms as TR
    from delayed_speech_synthesis_b200.units import HighGammaExtractor

    fs, n = 1000, 10
    x32 = O.synthetic_ecog(200, 128, fs, "ecog", session=4).astype(np.float32)
    sel = np.arange(0, 128, 2)
    ex = HighGammaExtractor(fs, 64, pre_transforms=[TR.ChannelSelection(sel)])
    ref = O.HighGammaExtractor(fs, 64, pre_transforms=[lambda d: O.select_columns(d, sel)])
    ex_cb = HighGammaExtractor(fs, 64, pre_transforms=[lambda d: d[:, sel]])      # unfoldable: the callable gets an ndarray
    for t in range(0, 200, n):
        pkt = O.pack_packet(x32[t:t + n].T)
        want = ref.extract_features(O.interpret_bytes(pkt))
        for e in (ex, ex_cb):
            got = e.extract_features_packet(pkt)
            assert isinstance(got, np.ndarray) and got.shape == want.shape
            assert np.abs(got - want).max() <= 1e-12
