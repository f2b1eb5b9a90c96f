"""Per-trial extraction of a recording as one batched launch (the part without file I/O):
`FeatureExtractionPipeline.get_features` (/root/reference/prepare_corpus.py:42-52) builds a fresh HighGammaExtractor per
trial and concatenates the features; `corpus.extract_trials` does the same with one segment table."""
from __future__ import annotations

import numpy as np
import pytest

from oracle import hga_oracle as O
from delayed_speech_synthesis_b200 import corpus

TRIALS = [("A", 100, 1300), ("SILENCE", 1320, 2100), ("B", 2080, 3490), ("C", 3500, 3520), ("D", 5000, 5990)]


def oracle_get_features(ecog, trials, fs, C, pre=None, post=None, tail=0.04):
    """prepare_corpus.py:42-52 written out with the oracle's extractor."""
    feats = []
    for _, start, stop in trials:
        ex = O.HighGammaExtractor(fs, C, pre_transforms=pre, post_transforms=post)
        feats.append(ex.extract_features(ecog[start:int(stop + (tail * fs)), :]))
    return np.concatenate(feats)


def test_trial_segments_follow_the_reference_slices():
    segs = corpus.trial_segments(TRIALS, 1000, 6000)
    assert segs == [(100, 1340), (1320, 2140), (2080, 3530), (3500, 3560), (5000, 6000)]   # last one clipped like a numpy slice
    assert corpus.trial_segments([("x", 10.0, 19.7)], 1000, 100, tail=0.0) == [(10, 19)]      # int() truncation of float indices
    assert corpus.trial_segments([("x", 7000, 8000)], 1000, 6000) == [(6000, 6000)]


@pytest.mark.gpu
def test_extract_trials_matches_per_trial_loop():
    fs, C = 1000, 128
    ecog = O.synthetic_ecog(6000, C, fs, "ecog", session=21)
    want = oracle_get_features(ecog, TRIALS, fs, C)
    got = corpus.extract_trials(ecog, TRIALS, fs, C)
    assert got.shape == want.shape and np.abs(got - want).max() <= 1e-9
    got32 = corpus.extract_trials(ecog, TRIALS, fs, C, precision="f32")
    assert np.abs(got32 - want).max() <= 1e-4


@pytest.mark.gpu
def test_extract_trials_short_trial_is_zero_padded_to_one_frame():
    fs, C = 1000, 32
    ecog = O.synthetic_ecog(3000, C, fs, "white", session=22)
    trials = [("A", 0, 500), ("tiny", 700, 705), ("B", 1000, 2000)]      # 45 rows < 50-row window -> frame-buffer case 2
    want = oracle_get_features(ecog, trials, fs, C)
    got = corpus.extract_trials(ecog, trials, fs, C)
    assert got.shape == want.shape and np.abs(got - want).max() <= 1e-9


@pytest.mark.gpu
def test_extract_trials_with_transform_chains():
    from delayed_speech_synthesis_b200 import transforms as TR

    fs = 1000
    ecog = O.synthetic_ecog(6000, 128, fs, "ecog", session=23)
    sel = np.arange(0, 128, 2)
    layout = np.arange(64) + 1
    grid = np.flip(layout.reshape(8, 8), axis=0)
    appl, comp = O.car_masks([19, 38], [grid], layout)
    patches = O.bad_channel_patches([19, 38], [grid], layout)
    pre_o = [lambda d: O.select_columns(d, sel), lambda d: O.common_average_reference(d, appl, comp)]
    post_o = [lambda f: O.bad_channel_correction(f, patches)]
    want = oracle_get_features(ecog, TRIALS, fs, 64, pre_o, post_o)
    pre = [TR.ChannelSelection(sel), TR.CommonAverageReferencing([19, 38], [grid], layout)]
    post = [TR.BadChannelCorrection([19, 38], [grid], layout)]
    got = corpus.extract_trials(ecog, TRIALS, fs, 64, pre, post)
    assert got.shape == want.shape and np.abs(got - want).max() <= 1e-9
    got_cb = corpus.extract_trials(ecog, TRIALS, fs, 64, pre_o, post_o)       # arbitrary callables: called per trial
    assert got_cb.shape == want.shape and np.abs(got_cb - want).max() <= 1e-9


# ------------------------------------------------------------------------------------------------ raw ingest (signal * gain)
def _raw_recording(T, C, dtype, seed):
    rng = np.random.Generator(np.random.Philox(key=[20241104, seed]))
    if np.issubdtype(dtype, np.integer):
        lim = min(np.iinfo(dtype).max, 30000)
        signal = rng.integers(-lim, lim, size=(T, C)).astype(dtype)
    else:
        signal = (rng.standard_normal((T, C)) * 3000.0).astype(dtype)
    gain = rng.uniform(0.02, 0.3, size=C)                 # uV per count, per channel (SourceChGain)
    return signal, gain


def test_mat_wrapper_reads_what_the_reference_wrapper_reads(tmp_path):
    """`BCI2000MatFile` (local/common.py:399-470) on a file written with the same field layout."""
    from scipy.io import savemat

    signal, gain = _raw_recording(200, 8, np.int16, 40)
    savemat(tmp_path / "KeywordReading_01.mat", {
        "signal": signal, "bad_channels": np.array(["chan3", "chan17"], dtype=object), "contaminated_electrodes": np.array([5, 6]),
        "parameters": {"SamplingRate": {"NumericValue": 1000}, "SourceChGain": {"NumericValue": gain}},
        "states": {"StimulusCode": np.repeat([0, 3, 0, 1, 0], 40).astype(np.uint8)}})
    f = corpus.BCI2000MatFile(str(tmp_path / "KeywordReading_01.mat"))
    assert f.fs == 1000 and f.bad_channels() == [3, 17] and f.contaminated_channels() == [5, 6]
    assert f.ordered_stimulus_codes() == [1, 3]
    raw, g = f.raw()
    assert raw.dtype == np.int16 and np.array_equal(raw, signal) and np.array_equal(g, gain)
    s = f.signals()
    assert s.dtype == np.float64 and np.array_equal(s, O.scaled_signals(signal, gain))
    assert np.array_equal(s, signal.astype(np.float64) * gain[None, :])


def test_raw_widening_is_lossless():
    for dt, want in ((np.uint8, np.int16), (np.int8, np.int16), (np.uint16, np.int32), (np.int64, np.float64), (np.float32, np.float32)):
        a = np.arange(6, dtype=dt).reshape(3, 2)
        b = corpus._raw_supported(a)
        assert b.dtype == want and np.array_equal(a, b)
    with pytest.raises(TypeError):
        corpus._raw_supported(np.zeros((2, 2), dtype=np.complex128))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.int16, np.int32, np.float32, np.float64])
def test_raw_ingest_is_bit_exact(dtype):
    import torch
    from delayed_speech_synthesis_b200 import transforms as TR

    signal, gain = _raw_recording(1031, 128, dtype, 41)
    want = O.scaled_signals(signal, gain)
    m = TR.ChannelMap(TR.ChannelMapSpec.identity(128))
    xd = torch.from_numpy(signal).cuda()
    assert np.array_equal(m.apply_raw(xd, gain).cpu().numpy(), want)
    assert np.array_equal(m.apply_raw(xd, 0.25).cpu().numpy(), O.scaled_signals(signal, 0.25))        # scalar gain
    assert np.array_equal(m.apply_raw(xd).cpu().numpy(), signal.astype(np.float64))                   # no gain
    sel = np.arange(127, -1, -2)
    ms = TR.ChannelMap(TR.fold([TR.ChannelSelection(sel)], 128))
    assert np.array_equal(ms.apply_raw(xd, gain).cpu().numpy(), want[:, sel])
    with pytest.raises(TypeError):
        m.apply_raw(xd.to(torch.int64), gain)


@pytest.mark.gpu
@pytest.mark.parametrize("C", [96, 1024, 6])
def test_raw_ingest_flat_stream_matches_rowwise(C):
    """The identity map over compact rows takes the flat k_rawscale stream (C % 4 == 0; 96 columns: 24 quads per row, so the
    grid is rounded to a multiple of 3 CTAs to keep each thread on fixed columns); padded rows and C = 6 take the row-wise
    k_chanmap.  Both must be `signal * gain` bit for bit (/root/reference/local/common.py:456-459)."""
    import torch
    from delayed_speech_synthesis_b200 import transforms as TR

    T = 4099
    signal, gain = _raw_recording(T, C, np.int16, 45)
    want = O.scaled_signals(signal, gain)
    m = TR.ChannelMap(TR.ChannelMapSpec.identity(C))
    xd = torch.from_numpy(signal).cuda()
    assert np.array_equal(m.apply_raw(xd, gain).cpu().numpy(), want)
    assert np.array_equal(m.apply_raw(xd).cpu().numpy(), signal.astype(np.float64))
    wide = torch.zeros((T, C + 4), dtype=torch.int16, device="cuda")
    wide[:, :C] = xd
    assert np.array_equal(m.apply_raw(wide[:, :C], gain).cpu().numpy(), want)
    for dt in (np.float32, np.float64, np.int32):
        sig, g = _raw_recording(517, C, dt, 46)
        assert np.array_equal(m.apply_raw(torch.from_numpy(sig).cuda(), g).cpu().numpy(), O.scaled_signals(sig, g))


@pytest.mark.gpu
def test_extract_trials_from_raw_counts():
    from delayed_speech_synthesis_b200 import transforms as TR

    fs, C = 1000, 128
    signal, gain = _raw_recording(6000, C, np.int16, 42)
    ecog = O.scaled_signals(signal, gain)
    want = oracle_get_features(ecog, TRIALS, fs, C)
    got = corpus.extract_trials(signal, TRIALS, fs, C, gain=gain)
    assert got.shape == want.shape and np.abs(got - want).max() <= 1e-9
    assert np.array_equal(got, corpus.extract_trials(ecog, TRIALS, fs, C))                            # same bits as the float64 upload
    sel = np.arange(0, 128, 2)
    want_s = oracle_get_features(ecog, TRIALS, fs, 64, pre=[lambda d: O.select_columns(d, sel)])
    got_s = corpus.extract_trials(signal, TRIALS, fs, 64, pre_transforms=[TR.ChannelSelection(sel)], gain=gain)
    assert np.abs(got_s - want_s).max() <= 1e-9
    got_cb = corpus.extract_trials(signal, TRIALS, fs, 64, pre_transforms=[lambda d: d[:, sel]], gain=gain)   # unfoldable chain
    assert np.abs(got_cb - want_s).max() <= 1e-9
    with pytest.raises(ValueError):
        corpus.extract_trials(signal, TRIALS, fs, C, gain=np.ones(7))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.int16, np.float32])
def test_offline_raw_host_slabs_match_float64_upload(dtype):
    """hga_extract_offline_raw_host == hga_extract_offline_host on signal * gain, bit for bit (same kernels, same rows),
    across several upload slabs, with a folded selection and the offline z-score."""
    from delayed_speech_synthesis_b200 import transforms as TR
    from delayed_speech_synthesis_b200.engine import HgaEngine

    fs, C = 1000, 128
    T = 400_000                                                      # > one 16 MiB float64 slab
    signal, gain = _raw_recording(T, C, dtype, 43)
    ecog = O.scaled_signals(signal, gain)
    mu, sigma = np.linspace(1.0, 2.0, C), np.linspace(0.5, 1.5, C)
    eng = HgaEngine(fs, C)
    ident = TR.ChannelMap(TR.ChannelMapSpec.identity(C), eng.device)
    segs = [(0, 150_000), (150_000, 150_060), (150_100, T)]
    want = eng.extract_offline(ecog, segs, mu, sigma)
    got = eng.extract_offline_raw(signal, ident, gain, segs, mu, sigma)
    assert got.shape == want.shape and np.array_equal(got, want)
    ref = O.HighGammaExtractor(fs, C, post_transforms=[O.ZScoreNormalization(mu, sigma)]).extract_features(ecog[:3000])
    assert np.abs(got[:len(ref)] - ref).max() <= 1e-9
    # strided host rows (a column block of a wider file) and a folded selection
    wide = np.zeros((T, C + 6), dtype=dtype)
    wide[:, :C] = signal
    assert np.array_equal(eng.extract_offline_raw(wide[:, :C], ident, gain, segs, mu, sigma), want)
    sel = np.arange(0, C, 2)
    eng64 = HgaEngine(fs, 64)
    m = TR.ChannelMap(TR.fold([TR.ChannelSelection(sel)], C), eng64.device)
    want_s = eng64.extract_offline(np.ascontiguousarray(ecog[:, sel]), segs)
    assert np.array_equal(eng64.extract_offline_raw(signal, m, gain, segs), want_s)
    with pytest.raises(ValueError):
        eng.extract_offline_raw(signal, m, gain, segs)               # map width != engine width
    with pytest.raises(TypeError):
        eng.extract_offline_raw(signal.astype(np.int64), ident, gain, segs)


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["exact", "f32"])
def test_offline_raw_host_other_precisions(precision):
    """The raw host call in the exact mode (single upload, sequential kernels) and in float32 mode: same bits as the float64
    host call of that mode on signal * gain."""
    from delayed_speech_synthesis_b200 import transforms as TR
    from delayed_speech_synthesis_b200.engine import HgaEngine

    fs, C, T = 1000, 64, 20_000
    signal, gain = _raw_recording(T, C, np.int32, 44)
    ecog = O.scaled_signals(signal, gain)
    eng = HgaEngine(fs, C, precision=precision)
    ident = TR.ChannelMap(TR.ChannelMapSpec.identity(C), eng.device)
    segs = [(0, 12_000), (12_000, T)]
    want = eng.extract_offline(ecog, segs)
    got = eng.extract_offline_raw(signal, ident, gain, segs)
    assert got.shape == want.shape and np.array_equal(got, want)
    assert np.array_equal(eng.extract_offline_raw(signal, ident, None, segs), eng.extract_offline(signal.astype(np.float64), segs))
