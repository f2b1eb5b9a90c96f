"""Corpus-scale offline extraction: sessions sharded over ranks (one process per GPU), baseline statistics combined with
ONE small collective (reference loops: prepare_corpus.py:179-236, baseline_offline.py:37-60)."""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Sequence


def shard_sessions(n_sessions: int, rank: int, world: int) -> List[int]:
    """Round-robin assignment of session indices to ranks (no data-path exchange: sessions are independent)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    return list(range(rank, n_sessions, world))


def shard_weighted(weights, rank: int, world: int) -> List[int]:
    """Session indices of `rank` when session s costs weights[s]: heaviest first, each to the least loaded rank (ties to the
    lowest rank, equal weights in index order) — the same table on every rank.  A day's baseline session is extracted twice
    (once raw for the statistics, once z-scored: prepare_corpus.py:190-199 then :209-227), so it weighs 2."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    load = [0.0] * world
    mine: List[int] = []
    for s in sorted(range(len(weights)), key=lambda i: (-float(weights[i]), i)):
        r = min(range(world), key=lambda k: (load[k], k))
        load[r] += float(weights[s])
        if r == rank:
            mine.append(s)
    return sorted(mine)


def gather_accumulators(acc, group=None):
    """All-gather a (k, 3, C) or (3, C) block of (n, mean, M2) accumulators -> (world * k, 3, C), rank-major.

    Works on CUDA tensors over NCCL (NVLink) and on CPU tensors over gloo (host-logic tests).  The merge itself is
    `ops.ChannelStats.merge` (a CUDA kernel, fixed order -> bit-identical mu/sigma on every rank)."""
    import torch
    import torch.distributed as dist

    a = acc.reshape(-1, 3, acc.shape[-1]).contiguous()
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return a
    world = dist.get_world_size(group)
    out = torch.empty((world * a.shape[0],) + tuple(a.shape[1:]), dtype=a.dtype, device=a.device)
    dist.all_gather_into_tensor(out, a, group=group)
    return out


def day_slots(days: Sequence[int]) -> Dict[int, int]:
    """Stable slot index per distinct day label (every rank must derive the same table)."""
    return {d: i for i, d in enumerate(sorted(set(days)))}


def merge_days(gathered, n_src: int, n_days: int):
    """(n_src, n_days, 3, C) gathered accumulators -> (mu, sigma), each (n_days, C): per (day, channel) a fixed-order Chan
    merge over the sources, then mu = mean, sigma = sqrt(M2 / n).  One kernel launch on CUDA tensors
    (`hga_stats_merge_days`); CPU tensors (gloo host-logic tests) take the same arithmetic in torch."""
    import torch

    g = gathered.reshape(n_src, n_days, 3, gathered.shape[-1])
    Cn = g.shape[-1]
    if g.is_cuda:
        from . import _lib
        from .engine import _stream_ptr

        g = g.contiguous()
        mu = torch.empty((n_days, Cn), dtype=torch.float64, device=g.device)
        sigma = torch.empty_like(mu)
        dev = g.device.index if g.device.index is not None else torch.cuda.current_device()
        with torch.cuda.device(dev):
            _lib.check(_lib.load().hga_stats_merge_days(dev, g.data_ptr(), n_src, n_days, Cn, mu.data_ptr(), sigma.data_ptr(), _stream_ptr(dev)))
        return mu, sigma
    n = torch.zeros((n_days, Cn), dtype=torch.float64)
    mean = torch.zeros_like(n)
    m2 = torch.zeros_like(n)
    for r in range(n_src):                                    # the kernel's chan_merge, source by source
        bn, bmean, bm2 = g[r, :, 0], g[r, :, 1], g[r, :, 2]
        tot = n + bn
        w = torch.where(tot > 0, bn / torch.where(tot > 0, tot, torch.ones_like(tot)), torch.zeros_like(tot))
        d = bmean - mean
        keep = tot > 0
        mean = torch.where(keep, mean + d * w, mean)
        m2 = torch.where(keep, m2 + bm2 + d * d * n * w, m2)
        n = torch.where(keep, tot, n)
    has = n > 0
    safe = torch.where(has, n, torch.ones_like(n))
    return torch.where(has, mean, torch.zeros_like(mean)), torch.where(has, torch.sqrt(m2 / safe), torch.ones_like(m2))


def _chan_merge(a, b):
    """(3, C) accumulators (n, mean, M2): Chan et al.'s pairwise merge, elementwise on the tensors' device."""
    import torch

    n = a[0] + b[0]
    safe = torch.where(n > 0, n, torch.ones_like(n))
    w, d = b[0] / safe, b[1] - a[1]
    return torch.stack([n, a[1] + d * w, a[2] + b[2] + d * d * a[0] * w])


def extract_corpus(n_sessions: int, day_of, baseline_sessions, get_session, extract, accumulate, *, n_channels: int, device=None,
                   group=None, on_features=None, timings: Optional[dict] = None, extract_stats=None):
    """The reference's corpus build (prepare_corpus.py:179-236) over the ranks of a process group, one process per GPU:

      1. `ZScoresFromSyllableRepetitions` (:183-199): per day, (n, mean, M2) of the features of that day's baseline
         session(s); every rank does this for the baseline sessions it owns (sessions are dealt by cost,
         `shard_weighted`: a baseline session is extracted twice), days it owns nothing of stay (0, 0, 0);
      2. ONE exchange: all days' accumulators in one all-gather (3 * C float64 per day and rank; NCCL over NVLink on CUDA
         tensors), then the same fixed-order merge on every rank (`merge_days`) -> bit-identical (mu, sigma) everywhere;
      3. the main loop (:209-234): every owned session extracted and z-scored with its day's (mu, sigma) (:226-227).

    day_of[s]: day label of session s.  baseline_sessions: session ids whose features define their day's statistics.
    get_session(s) -> the session's recording (device tensor).  extract(x, mu, sigma) -> (F, C) features (mu = None: raw
    log-power).  accumulate(features) -> (3, C) tensor (n, mean, M2).  on_features(s, features): what the caller does with
    a finished session (the reference writes an HDF5 file, :233-234).  timings: optional dict that receives CUDA events
    bracketing the exchange (`exchange_events`) when the tensors live on a device.  extract_stats(x) -> (3, C): when given,
    it replaces accumulate(extract(x, None, None)) for the baseline sessions (`HgaEngine.extract_offline_stats`: the
    statistics leave the extraction kernels, the features are not read back).
    Returns (mu, sigma, slots): (n_days, C) tensors and the day -> row table."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank(group) if world > 1 else 0
    slots = day_slots([day_of[s] for s in range(n_sessions)])
    n_days = len(slots)
    base = set(int(b) for b in baseline_sessions)
    mine = shard_weighted([2 if s in base else 1 for s in range(n_sessions)], rank, world)
    acc = None
    for s in mine:
        if s in base:
            a = extract_stats(get_session(s)) if extract_stats is not None else accumulate(extract(get_session(s), None, None))
            if acc is None:
                acc = torch.zeros((n_days, 3, n_channels), dtype=torch.float64, device=a.device)
            d = slots[day_of[s]]
            acc[d] = _chan_merge(acc[d], a)          # several baseline sessions of one day on this rank: in session order
    if acc is None:
        dev = torch.device("cuda", device) if device is not None else torch.device("cpu")
        acc = torch.zeros((n_days, 3, n_channels), dtype=torch.float64, device=dev)
    ev = None
    if acc.is_cuda and timings is not None:
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record()
    gathered = gather_accumulators(acc, group)                      # (world * n_days, 3, C), rank-major
    mu, sigma = merge_days(gathered, world, n_days)
    if ev is not None:
        ev[1].record()
        timings.setdefault("exchange_events", []).append(ev)
    for s in mine:
        d = slots[day_of[s]]
        f = extract(get_session(s), mu[d], sigma[d])
        if on_features is not None:
            on_features(s, f)
    return mu, sigma, slots


# ---------------------------------------------------------------------------------------------- per-trial extraction
def trial_segments(trial_indices, fs: float, n_rows: int, tail: float = 0.04):
    """The row ranges `FeatureExtractionPipeline.get_features` slices out of a recording (prepare_corpus.py:47-50):
    `ecog[start:int(stop + tail * fs), :]` per (label, start, stop) trial, clipped to the recording like a numpy slice
    (trial indices are sample positions, i.e. non-negative)."""
    segs = []
    for _, start, stop in trial_indices:
        a = min(max(int(start), 0), n_rows)
        b = min(max(int(stop + (tail * fs)), 0), n_rows)
        segs.append((a, max(a, b)))
    return segs


def extract_trials(ecog, trial_indices, fs: int, nb_electrodes: int, pre_transforms=None, post_transforms=None,
                   window_length: float = 0.05, window_shift: float = 0.01, l_freq: float = 70, h_freq: float = 170,
                   tail: float = 0.04, precision: str = "f64", device: Optional[int] = None, gain=None):
    """`FeatureExtractionPipeline.get_features` (prepare_corpus.py:42-52) as ONE batched launch (the part
    that needs no file I/O): every trial is filtered by a fresh extractor (initial state sosfilt_zi, fresh frame buffer),
    the features of all trials are concatenated.  The reference builds a `HighGammaExtractor` per trial and loops in Python.

    ecog: (T, C_in) float64 ndarray, e.g. `BCI2000MatFile.signals()`.  pre/post transforms as for `HighGammaExtractor`:
    chains of the reference's selection / CAR / bad-channel objects run as one channel-map kernel over the whole recording
    (they are row-wise, so slicing before or after is the same); anything else is called per trial like the reference does.
    gain: when given, `ecog` is the file's raw `signal` matrix (int16 / int32 / float32 / float64 counts) and `gain` its
    `SourceChGain` vector (`BCI2000MatFile.raw()`): the product `signal * gain` of `BCI2000MatFile.signals()`
    (local/common.py:456-459) is formed on the device in the load stage of the channel-map kernel, so the upload is the raw
    element size (2 or 4 B per channel-sample instead of 8).
    Returns a (sum of frames, nb_electrodes) float64 ndarray."""
    import numpy as np
    import torch

    from . import transforms
    from .engine import HgaEngine

    ecog = np.asarray(ecog)
    if ecog.ndim != 2:
        raise ValueError(f"expected (T, C) data, got {ecog.shape}")
    if gain is not None:
        ecog = _raw_supported(ecog)
        gain = np.asarray(gain, dtype=np.float64)
        if gain.ndim > 1 or (gain.ndim == 1 and len(gain) not in (1, ecog.shape[1])):
            raise ValueError(f"gain of shape {gain.shape} does not broadcast over {ecog.shape[1]} channels")
    elif ecog.dtype != np.float64:
        ecog = ecog.astype(np.float64)
    eng = HgaEngine(fs, nb_electrodes, window_length, window_shift, l_freq, h_freq, precision=precision, device=device)
    dev = eng.device
    segs = trial_segments(trial_indices, fs, len(ecog), tail)
    L = eng.frame_len
    spec = transforms.fold(pre_transforms, ecog.shape[1]) if pre_transforms else None
    if pre_transforms and (spec is None or spec.c_out != nb_electrodes):
        # unfoldable chain: the user's callables see each trial slice, exactly as the reference calls them; the results are
        # laid end to end and become the segment table of the one launch
        scaled = (lambda v: v) if gain is None else (lambda v: v * gain)   # what signals() hands the user's callables
        parts = [np.asarray(_compose(pre_transforms)(scaled(ecog[a:b]))) for a, b in segs]
        offs = np.concatenate([[0], np.cumsum([len(p) for p in parts])])
        x = torch.from_numpy(np.ascontiguousarray(np.concatenate(parts) if parts else np.zeros((0, nb_electrodes)), dtype=np.float64)).cuda(dev)
        segs = [(int(offs[i]), int(offs[i + 1])) for i in range(len(parts))]
    else:
        if gain is not None and not post_transforms and segs and all(b - a >= L for a, b in segs):
            # every trial fills a window: upload slabs of raw counts, expand and extract on the device, all inside one call
            cmap = transforms.ChannelMap(spec or transforms.ChannelMapSpec.identity(ecog.shape[1]), dev)
            return eng.extract_offline_raw(ecog, cmap, gain, segs)
        x = torch.from_numpy(np.ascontiguousarray(ecog)).cuda(dev)
        if gain is None and spec is not None and not post_transforms and segs and all(b - a >= L for a, b in segs):
            # the folded chain runs inside the extraction kernel's load stage (one launch, no mapped recording in memory)
            return eng.extract_offline_mapped(x, transforms.ChannelMap(spec, dev), segs).cpu().numpy()
        if gain is not None:
            x = transforms.ChannelMap(spec or transforms.ChannelMapSpec.identity(ecog.shape[1]), dev).apply_raw(x, gain)
        elif spec is not None:
            x = transforms.ChannelMap(spec, dev).apply(x)
    if x.shape[1] != nb_electrodes:
        raise ValueError(f"expected {nb_electrodes} channels after the pre-transforms, got {x.shape[1]}")
    long_ = [(i, s) for i, s in enumerate(segs) if s[1] - s[0] >= L]
    feats = [None] * len(segs)
    if long_:
        out = eng.extract_offline(x, [s for _, s in long_])
        rows = np.concatenate([[0], np.cumsum([eng.num_frames(s[1] - s[0]) for _, s in long_])])
        for k, (i, _) in enumerate(long_):
            feats[i] = out[int(rows[k]):int(rows[k + 1])]
    for i, (a, b) in enumerate(segs):
        if feats[i] is None:  # shorter than one window: a fresh extractor zero-pads it to one frame (frame-buffer case 2)
            if b - a < 1:
                feats[i] = x.new_zeros((0, nb_electrodes))
            else:
                eng.reset()
                feats[i] = eng.push(x[a:b].contiguous())
    res = torch.cat(feats) if feats else x.new_zeros((0, nb_electrodes))
    if post_transforms:
        pspec = transforms.fold(post_transforms, nb_electrodes)
        if pspec is not None and pspec.c_out == nb_electrodes and len(res):
            res = transforms.ChannelMap(pspec, dev).apply(res.contiguous())
        else:
            # the reference applies post-transforms per trial; row-wise callables commute with the concatenation
            return np.concatenate([np.asarray(_compose(post_transforms)(f.cpu().numpy())) for f in feats]) if feats else res.cpu().numpy()
    return res.cpu().numpy()


def _raw_supported(a):
    """Element types the raw-ingest kernel reads (float64, float32, int32, int16); narrower or unsigned integers are widened
    losslessly on the host, 64-bit integers go to float64 as numpy's own promotion with a float64 gain would."""
    import numpy as np

    if a.dtype in (np.float64, np.float32, np.int32, np.int16):
        return a
    if a.dtype in (np.int8, np.uint8):
        return a.astype(np.int16)
    if a.dtype == np.uint16:
        return a.astype(np.int32)
    if a.dtype.kind in "iub" or a.dtype == np.float16:
        return a.astype(np.float64)
    raise TypeError(f"raw signal of dtype {a.dtype} is not real-valued")


class BCI2000MatFile:
    """The accessors of the reference's `.mat` wrapper that feed this path (local/common.py:399-470): sampling rate, channel
    bookkeeping and the signal matrix.  `raw()` returns the two factors of `signals()` so that `extract_trials(..., gain=)`
    can form the product on the device.  `trial_indices` / `stimuli_indices` (local/common.py:432-456) come from the
    vectorised state-vector scans of `trials.py`; their (label, start, stop) list is what `extract_trials` takes."""

    def __init__(self, mat_filename: str):
        from scipy.io import loadmat

        self.mat_filename = mat_filename
        self.mat = loadmat(self.mat_filename, simplify_cells=True)
        self.fs = self.mat['parameters']['SamplingRate']['NumericValue']

    def bad_channels(self):
        if 'bad_channels' not in self.mat:
            return None
        bad = self.mat['bad_channels']
        if hasattr(bad, "tolist"):
            bad = bad.tolist()
        bad = bad if isinstance(bad, list) else [bad]
        return [int(name[4:]) for name in bad]          # 'chan17' -> 17

    def contaminated_channels(self):
        if 'contaminated_electrodes' not in self.mat:
            return None
        c = self.mat['contaminated_electrodes']
        return [c] if isinstance(c, int) else c.tolist()

    def raw(self):
        return self.mat['signal'], self.mat['parameters']['SourceChGain']['NumericValue']

    def signals(self):
        signal, gain = self.raw()
        return signal * gain

    def trial_indices(self, min_trial_length=None):
        from . import trials
        return trials.trial_indices(self.mat, self.mat_filename, self.fs, min_trial_length)

    def stimuli_indices(self):
        from . import trials
        return trials.stimuli_indices(self.mat, self.mat_filename)

    def ordered_stimulus_codes(self):
        import numpy as np

        return sorted(np.unique(self.mat['states']['StimulusCode']).tolist())[1:]


def _compose(functions):
    """`HighGammaExtractor._compose_functions` (local/units.py:134-136)."""
    from functools import reduce
    return reduce(lambda f, g: lambda v: g(f(v)), functions, lambda v: v)
