"""Trial / stimulus index extraction from a BCI2000 `StimulusCode` state vector, and the segment table the batched
extraction takes.

Mirrors, with the same names and results, the reference's `Experiment` family (local/common.py:676-880):
`Experiment._determine_trial_boundaries` (:693-695), `SyllableRepetition.get_trial_indices / get_stimuli_indices`
(:745-812), `KeywordReading` (:815-844), `ExperimentMapping` (:847-880), and `BCI2000MatFile.trial_indices /
stimuli_indices` (:432-454).  The reference walks the state vector sample by sample in Python (`_stimuli_extraction`,
:793-807, one interpreter iteration per sample: 3.6 M for an hour at 1 kHz); here every list comes from run-length
boundaries found with vectorised NumPy comparisons.  Host-side bookkeeping: no arithmetic of the hot path lives here,
its product is the `(start, length)` table of `hga_extract_offline`.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple, Union

import numpy as np

TrialIndices = Tuple[str, int, int]


def _runs(code: np.ndarray) -> np.ndarray:
    """Positions where the state vector changes value: `np.where(c[:-1] != c[1:])[0] + 1` (local/common.py:694)."""
    code = np.asarray(code)
    return np.flatnonzero(code[:-1] != code[1:]) + 1


class Experiment:
    """Interface of local/common.py:676-741: `get_trial_indices`, `get_stimuli_indices`, baseline windows, run indices."""

    def __init__(self, stimulus_code: np.ndarray, stimuli: Union[Dict[int, str], List[str]]):
        self.stimulus_code = np.asarray(stimulus_code)
        self.stimuli = stimuli
        if isinstance(self.stimuli, list):  # position in the list + 1 is the code; 0 is "no stimulus" (:684-686)
            self.stimuli = {(index + 1): item for index, item in enumerate(self.stimuli)}

    def __repr__(self):
        return (f'{self.__class__.__name__}(len: {len(self.stimulus_code)} samples, with {len(self.stimuli)} stimuli '
                f'across {len(self.get_trial_indices())} trials)')

    def _determine_trial_boundaries(self) -> List[Tuple[int, int]]:
        diff = _runs(self.stimulus_code)
        return list(zip(diff[::].tolist(), diff[1::].tolist()))

    @staticmethod
    def trial_indices_to_lab(filename: str, trial_indices: Sequence[TrialIndices], fs: int):
        with open(filename, 'w') as f:
            for label, start, stop in trial_indices:
                f.write(f'{start / fs:.03f}\t{stop / fs:.03f}\t{label}\n')

    def get_trial_indices(self) -> List[TrialIndices]:
        raise NotImplementedError

    def get_stimuli_indices(self) -> List[TrialIndices]:
        raise NotImplementedError

    def get_webfm_baseline_windows(self, fs: int, length: float = 0.8) -> List[TrialIndices]:
        return [('BL', int(start - length * fs), start) for _, start, _ in self.get_stimuli_indices()]

    def get_experiment_run_indices(self) -> TrialIndices:
        start = self._determine_trial_boundaries()[0][0]
        stop = self.get_trial_indices()[-1][2]
        return 'Experiment run', start, stop

    @staticmethod
    def get_stimuli_values() -> Optional[list]:
        return None

    # ------------------------------------------------------------------ shared by both experiments (:793-807, :821-835)
    def _nonzero_runs(self) -> List[TrialIndices]:
        """Segments with StimulusCode != 0, closed by the first 0 after them (a run still open at the end of the
        recording is dropped, as the reference's sample loop drops it; a change between two non-zero codes does not close a
        segment: the label stays that of the run's first sample)."""
        c = self.stimulus_code
        if len(c) == 0:
            return []
        nz = (c != 0).astype(np.int8)
        edges = np.diff(np.concatenate([[0], nz]))          # +1 at the first sample of a non-zero run, -1 at the zero closing it
        starts = np.flatnonzero(edges == 1)
        stops = np.flatnonzero(edges == -1)
        return [(self.stimuli[c[a].item()], int(a), int(b)) for a, b in zip(starts[:len(stops)], stops)]


class SyllableRepetition(Experiment):
    """Syllables are presented acoustically (code != 0), then the patient repeats them (code == 0) (:744-812)."""

    def get_trial_indices(self) -> List[TrialIndices]:
        c = self.stimulus_code
        trials = self._determine_trial_boundaries()
        start, stop = trials[1]                              # trial length = length of the second segment (:765-768)
        trial_length = min(stop - start, len(c))
        trials.append((trials[-1][1], trials[-1][1] + trial_length))   # the last trial has no closing change (:781-783)
        # codes move from the presentation segments (even positions) to the speaking segments that follow them, the
        # presentation segments are zeroed (:747-762); a trial is kept when its first sample then carries a code
        out = []
        for k, (a, b) in enumerate(trials):
            if k % 2 == 1:
                code = c[trials[k - 1][0]]
                if a < len(c) and b > a and code != 0:       # an empty or out-of-range slice assignment changes nothing
                    out.append((self.stimuli[code.item()], a, b))
            # even positions were zeroed unless the zeroing slice was empty / outside the vector; then stim_codes[start]
            # is whatever a later speaking segment wrote there or the original code
            else:
                if a < len(c) and not (b > a):
                    code = self._code_after_swap(trials, a)
                    if code != 0:
                        out.append((self.stimuli[int(code)], a, b))
        return out

    def _code_after_swap(self, trials, pos) -> int:
        """Value of the swapped code vector at `pos` (only needed for degenerate, empty segments)."""
        c = self.stimulus_code
        new = c.copy()
        for k, (a, b) in enumerate(trials[1::2]):
            new[a:b] = c[trials[::2][k][0]]
        for a, b in trials[::2]:
            new[a:b] = 0
        return int(new[pos])

    def get_stimuli_indices(self) -> List[TrialIndices]:
        return self._nonzero_runs()


class KeywordReading(Experiment):
    """Keywords shown on screen; trial and stimulus indices coincide (:815-844)."""

    def get_trial_indices(self) -> List[TrialIndices]:
        return self.get_stimuli_indices()

    def get_stimuli_indices(self) -> List[TrialIndices]:
        return self._nonzero_runs()


class ExperimentMapping(dict):
    """Experiment name in the file name -> class (:847-880)."""

    def __init__(self):
        super().__init__()
        self.update({'SyllableRepetition': SyllableRepetition, 'KeywordReading': KeywordReading, 'KeywordSynthesis': KeywordReading})

    @staticmethod
    def get_experiment_class(mat_filename: str):
        filename = os.path.basename(mat_filename)
        mapping = ExperimentMapping()
        for key in mapping.keys():
            if key in filename:
                return mapping[key]
        return None

    @staticmethod
    def extract_stimuli_values(mat: dict) -> List[str]:
        stimuli = mat['parameters']['Stimuli']['Value']
        if stimuli.ndim == 1:
            return [stimuli[0]]
        return stimuli[0].tolist()


def trial_indices(mat: dict, mat_filename: str, fs: float, min_trial_length: Optional[float] = None) -> List[TrialIndices]:
    """`BCI2000MatFile.trial_indices` (local/common.py:432-445) on a loaded `.mat` dictionary."""
    stimuli = ExperimentMapping.extract_stimuli_values(mat)
    experiment = ExperimentMapping.get_experiment_class(mat_filename)(mat['states']['StimulusCode'], stimuli)
    trials = experiment.get_trial_indices()
    if min_trial_length is not None:
        nb_min_samples = min_trial_length * fs
        trials = [(label, start, max(stop, start + nb_min_samples)) for label, start, stop in trials]
    return trials


def stimuli_indices(mat: dict, mat_filename: str) -> List[TrialIndices]:
    """`BCI2000MatFile.stimuli_indices` (local/common.py:447-456)."""
    stimuli = ExperimentMapping.extract_stimuli_values(mat)
    return ExperimentMapping.get_experiment_class(mat_filename)(mat['states']['StimulusCode'], stimuli).get_stimuli_indices()


def segment_table(trials: Sequence[TrialIndices], fs: float, n_rows: int, frame_len: int, hop: int, tail: float = 0.04):
    """The trials of a recording as the arrays `hga_extract_offline` takes: (seg_start, seg_len, out_row, labels).

    Row ranges are the reference's slices `ecog[start:int(stop + tail * fs)]` (prepare_corpus.py:47-50) clipped to the
    recording; `out_row[i]` is where trial i's frames start when all trials are laid end to end (`np.vstack`, :52); trials
    shorter than one window get `seg_len < frame_len` and zero frames here (the frame buffer's zero-padding path handles them)."""
    start = np.empty(len(trials), np.int64)
    length = np.empty(len(trials), np.int64)
    for i, (_, a, b) in enumerate(trials):
        a = min(max(int(a), 0), n_rows)
        b = min(max(int(b + tail * fs), 0), n_rows)
        start[i], length[i] = a, max(0, b - a)
    frames = np.where(length >= frame_len, (length - frame_len) // hop + 1, 0)
    out_row = np.concatenate([[0], np.cumsum(frames)[:-1]]).astype(np.int64) if len(trials) else np.zeros(0, np.int64)
    return start, length, out_row, [t[0] for t in trials]
