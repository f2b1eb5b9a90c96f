"""Trial / stimulus index extraction (delayed_speech_synthesis_b200/trials.py) against tests/golden/trials.npz: the
reference's own Experiment classes (local/common.py:676-880) executed on synthetic StimulusCode vectors by
tests/golden/make_golden.py.  Index work: every list must match exactly."""
from pathlib import Path

import numpy as np
import pytest

from delayed_speech_synthesis_b200 import trials as TR

GOLD = np.load(Path(__file__).parent / "golden" / "trials.npz")
CASES = sorted({k.split(".")[0] for k in GOLD.files})


def _experiment(name):
    paradigm = str(GOLD[f"{name}.paradigm"])
    stimuli = [f"stim{j + 1:02d}" for j in range(int(GOLD[f"{name}.n_codes"]))]
    cls = TR.ExperimentMapping.get_experiment_class(f"/data/day1/{paradigm}_Run01.mat")
    return cls(GOLD[f"{name}.code"], stimuli), stimuli, paradigm


def _check(name, kind, got):
    assert [g[0] for g in got] == [str(s) for s in GOLD[f"{name}.{kind}.labels"]]
    want = GOLD[f"{name}.{kind}.bounds"]
    assert np.array_equal(np.array([[g[1], g[2]] for g in got], dtype=np.int64).reshape(-1, 2), want)


@pytest.mark.parametrize("name", CASES)
def test_trial_and_stimulus_indices_match_the_reference(name):
    exp, _, _ = _experiment(name)
    _check(name, "trials", exp.get_trial_indices())
    _check(name, "stimuli", exp.get_stimuli_indices())
    bl = exp.get_webfm_baseline_windows(1000)
    assert np.array_equal(np.array([[b[1], b[2]] for b in bl], dtype=np.int64).reshape(-1, 2), GOLD[f"{name}.baseline.bounds"])
    assert all(b[0] == "BL" for b in bl)
    if f"{name}.run" in GOLD.files:
        run = exp.get_experiment_run_indices()
        assert run[0] == "Experiment run" and [run[1], run[2]] == GOLD[f"{name}.run"].tolist()
    else:   # the reference raises on a recording without any stimulus; so does the mirror
        with pytest.raises(IndexError):
            exp.get_experiment_run_indices()


def test_mapping_and_mat_file_entry_points():
    assert TR.ExperimentMapping.get_experiment_class("x/KeywordSynthesis_03.mat") is TR.KeywordReading
    assert TR.ExperimentMapping.get_experiment_class("x/SyllableRepetition_R02.mat") is TR.SyllableRepetition
    assert TR.ExperimentMapping.get_experiment_class("x/SomethingElse.mat") is None
    name = "key_jitter"
    exp, stimuli, paradigm = _experiment(name)
    mat = {"parameters": {"Stimuli": {"Value": np.array([stimuli, ["a"] * len(stimuli)], dtype=object)}}, "states": {"StimulusCode": GOLD[f"{name}.code"]}}
    got = TR.trial_indices(mat, f"{paradigm}_Run01.mat", fs=1000)
    _check(name, "trials", got)
    _check(name, "stimuli", TR.stimuli_indices(mat, f"{paradigm}_Run01.mat"))
    # BCI2000MatFile.trial_indices(min_trial_length) (local/common.py:441-443): stop = max(stop, start + min_len * fs)
    clamped = TR.trial_indices(mat, f"{paradigm}_Run01.mat", fs=1000, min_trial_length=1.8)
    for (l0, a0, b0), (l1, a1, b1) in zip(got, clamped):
        assert l0 == l1 and a0 == a1 and b1 == max(b0, a0 + 1.8 * 1000)
    one = {"parameters": {"Stimuli": {"Value": np.array(["only"], dtype=object)}}, "states": {"StimulusCode": np.zeros(4)}}
    assert TR.ExperimentMapping.extract_stimuli_values(one) == ["only"]


def test_lab_file(tmp_path):
    exp, _, _ = _experiment("key_single")
    p = tmp_path / "t.lab"
    TR.Experiment.trial_indices_to_lab(str(p), exp.get_trial_indices(), 1000)
    assert p.read_text() == "0.300\t0.800\tstim02\n"


def test_segment_table_is_the_reference_slicing():
    """prepare_corpus.py:47-52: features of ecog[start:int(stop + 0.04 * fs)] per trial, stacked."""
    exp, _, _ = _experiment("key_open_end")
    tr = exp.get_trial_indices()
    n_rows, fs, L, hop = len(GOLD["key_open_end.code"]), 1000, 50, 10
    start, length, out_row, labels = TR.segment_table(tr, fs, n_rows, L, hop)
    x = np.arange(n_rows)
    row = 0
    for i, (label, a, b) in enumerate(tr):
        sl = x[a:int(b + 0.04 * fs)]
        assert start[i] == (sl[0] if len(sl) else min(a, n_rows)) and length[i] == len(sl) and labels[i] == label
        assert out_row[i] == row
        row += (len(sl) - L) // hop + 1 if len(sl) >= L else 0
    # a trial shorter than one window holds no frame and takes no output row
    s2, l2, o2, _ = TR.segment_table([("a", 0, 5), ("b", 100, 400)], fs, 1000, L, hop)
    assert l2.tolist() == [45, 340] and o2.tolist() == [0, 0]
    # clipped at the end of the recording
    s3, l3, _, _ = TR.segment_table([("a", 900, 990)], fs, 1000, L, hop)
    assert (s3[0], l3[0]) == (900, 100)
    assert TR.segment_table([], fs, 1000, L, hop)[0].shape == (0,)


def test_mat_wrapper_trial_indices(tmp_path):
    """`BCI2000MatFile.trial_indices / stimuli_indices` (local/common.py:432-456) on a file with the reference's field layout."""
    from scipy.io import savemat

    from delayed_speech_synthesis_b200 import corpus

    name = "key_regular"
    code = GOLD[f"{name}.code"]
    stimuli = [f"stim{j + 1:02d}" for j in range(int(GOLD[f"{name}.n_codes"]))]
    cell = np.empty((2, len(stimuli)), dtype=object)
    cell[0, :] = stimuli
    cell[1, :] = "x"
    path = tmp_path / "KeywordReading_Run07.mat"
    savemat(path, {"signal": np.zeros((len(code), 2), np.int16),
                   "parameters": {"SamplingRate": {"NumericValue": 1000}, "SourceChGain": {"NumericValue": np.ones(2)}, "Stimuli": {"Value": cell}},
                   "states": {"StimulusCode": code.astype(np.uint8)}})
    f = corpus.BCI2000MatFile(str(path))
    _check(name, "trials", f.trial_indices())
    _check(name, "stimuli", f.stimuli_indices())
    long = f.trial_indices(min_trial_length=2.5)
    assert all(b - a >= 2500 for _, a, b in long)
