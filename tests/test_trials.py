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
