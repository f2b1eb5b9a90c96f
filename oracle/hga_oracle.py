"""CPU ORACLE for the HGA front-end hot path — TEST INFRASTRUCTURE ONLY.

This module is a NumPy/SciPy restatement of the reference's high-gamma-activity (HGA) feature path.  Only `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of `bench.py` may import it; the product
(`delayed_speech_synthesis_b200`) never does and fails loudly when its CUDA library is missing.

Parity status: the reference ships NO tests, golden vectors or fixtures for this path, so the
oracle is pinned against the reference *executed here* instead:
  * `compute_log_power_features`  == the compiled reference extension (oracle/_ref, built by oracle/build_ref.py from
    /root/reference/extensions/hga/hga_optimized.pyx) — checked in tests/test_oracle.py and frozen in tests/golden/.
  * the IIR arithmetic lives in third-party SciPy (`scipy.signal.sosfilt`, pinned scipy==1.11.4 in the reference's
    requirements.txt:2; 1.18.1 installed here).  `sosfilt_df2t` below restates its published DF2T recurrence and is
    asserted bit-identical to `scipy.signal.sosfilt` in tests/test_oracle.py.
  * filter design goes through third-party `mne.filter.create_filter` (mne==1.6.1, requirements.txt:11; NOT installed
    here).  Its IIR branch reduces to `scipy.signal.iirfilter(N=order, Wn, btype, ftype='butter', output='sos')`;
    that step is "parity unpinned" (no reference fixture holds the coefficients) — the SOS matrices are an *input* to
    every kernel, so a design-side discrepancy cannot hide in the CUDA path.

Every function cites the reference file:line it follows (paths relative to /root/reference).
"""
from __future__ import annotations

from functools import reduce
from typing import Callable, Iterable, List, Optional, Sequence, Tuple

import numpy as np
from scipy.signal import iirfilter, sosfilt, sosfilt_zi

EPS_LOG = 0.01  # extensions/hga/hga_optimized.pyx:46

# The reference calls libm `log` (hga_optimized.pyx:1,46).  `np.log` dispatches to a SIMD kernel that may differ by
# 1 ulp, so the oracle goes through `math.log` (libm) element-wise to stay bit-faithful.
import math as _math
_libm_log = np.frompyfunc(_math.log, 1, 1)


def _log(a: np.ndarray) -> np.ndarray:
    return _libm_log(a).astype(np.float64)


# ----------------------------------------------------------------------------------------------------------------------
# Filter design and initial state                                                     local/units.py:122-132, 138-143
# ----------------------------------------------------------------------------------------------------------------------
def create_filter_sos(fs: float, l_freq: float, h_freq: float, order: int = 8) -> np.ndarray:
    """SOS Butterworth design as `HighGammaExtractor.create_filter(...)["sos"]` yields it (local/units.py:138-143).

    mne 1.6.1 `create_filter(None, sr, l, h, ..., method='iir', iir_params={'order': 8, 'ftype': 'butter'})`:
    `l < h` -> band-pass with Wn = [l, h] / (fs/2); `l > h` -> band-stop with Wn = [h, l] / (fs/2)
    (`construct_iir_filter` forwards `order` as `N` to `scipy.signal.iirfilter(..., output='sos')`).
    """
    nyq = fs / 2.0
    if l_freq < h_freq:
        return iirfilter(order, np.array([l_freq, h_freq]) / nyq, btype="bandpass", ftype="butter", output="sos")
    return iirfilter(order, np.array([h_freq, l_freq]) / nyq, btype="bandstop", ftype="butter", output="sos")


def design_hga_filters(fs: float, l_freq: float = 70, h_freq: float = 170) -> Tuple[np.ndarray, np.ndarray]:
    """(hg_filter, fh_filter): 70-170 Hz band-pass and the 118-122 Hz line-harmonic notch (local/units.py:124-126)."""
    return create_filter_sos(fs, l_freq, h_freq), create_filter_sos(fs, 122, 118)


def tiled_zi(sos: np.ndarray, nb_electrodes: int) -> np.ndarray:
    """`sosfilt_zi(sos)` repeated per electrode to (n_sections, 2, C), NOT scaled by x[0] (local/units.py:128-132)."""
    zi = sosfilt_zi(sos)
    return np.repeat(zi, nb_electrodes, axis=-1).reshape([zi.shape[0], zi.shape[1], -1])


def sosfilt_df2t(sos: np.ndarray, x: np.ndarray, zi: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Scalar restatement of SciPy's `_sosfilt` (transposed direct form II), time along axis 0.

    Per channel, sample, section:
        y  = b0*x + z0 ;  z0' = (b1*x - a1*y) + z1 ;  z1' = b2*x - a2*y ;  x <- y
    Pure-Python float loop (no FMA contraction) — use only on small inputs.  x: (T, C); zi: (S, 2, C).
    """
    x = np.asarray(x, dtype=np.float64)
    T, C = x.shape
    y = np.empty_like(x)
    z = np.array(zi, dtype=np.float64, copy=True)
    coef = [tuple(float(v) for v in row) for row in np.asarray(sos, dtype=np.float64)]
    for c in range(C):
        zc = [[float(z[s, 0, c]), float(z[s, 1, c])] for s in range(len(coef))]
        for t in range(T):
            u = float(x[t, c])
            for s, (b0, b1, b2, _a0, a1, a2) in enumerate(coef):
                z0, z1 = zc[s]
                v = b0 * u + z0
                zc[s][0] = (b1 * u - a1 * v) + z1
                zc[s][1] = b2 * u - a2 * v
                u = v
            y[t, c] = u
        for s in range(len(coef)):
            z[s, 0, c], z[s, 1, c] = zc[s]
    return y, z


# ----------------------------------------------------------------------------------------------------------------------
# Framing                                                                       extensions/hga/hga_optimized.pyx:8-47
# ----------------------------------------------------------------------------------------------------------------------
def window_indices(rows: int, sr: int, window_length: float, window_shift: float) -> Tuple[np.ndarray, np.ndarray]:
    """(start, stop) row indices of every window, with the reference's C-`float` arithmetic.

    hga_optimized.pyx:27 declares `float window_length, float window_shift` and `int sr`, so
      :36  num_windows = <int> floor((rows - wl*sr) / (ws*sr)) + 1        -- all float32 until libm `floor(double)`
      :43  start = int(round((win*ws)*sr))                               -- float32 products, libm `round(double)`
      :44  stop  = int(round(start + wl*sr))                             -- int + float32 -> float32
    """
    f32 = np.float32
    wl, ws, srf = f32(window_length), f32(window_shift), f32(sr)
    num = int(np.floor(np.float64((f32(rows) - wl * srf) / (ws * srf)))) + 1
    if num <= 0:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    win = np.arange(num, dtype=np.int64).astype(f32)
    c_round = lambda v: np.copysign(np.floor(np.abs(v) + 0.5), v)  # libm round(): half away from zero
    start = c_round(((win * ws) * srf).astype(np.float64)).astype(np.int64)
    stop = c_round((start.astype(f32) + wl * srf).astype(np.float64)).astype(np.int64)
    return start, stop


def compute_log_power_features(data: np.ndarray, sr: int, window_length: float, window_shift: float) -> np.ndarray:
    """log(mean(x^2 over window) + 0.01) per window and channel (hga_optimized.pyx:8-22, 27-47).

    The reference sums sequentially in float64; `np.add.reduce` over <=100 rows is pairwise-free only below 8
    elements, so the sum here is done with an explicit sequential accumulation to stay bit-faithful.
    """
    data = np.asarray(data, dtype=np.float64)
    start, stop = window_indices(data.shape[0], sr, window_length, window_shift)
    out = np.empty((len(start), data.shape[1]), dtype=np.float64)
    sq = data * data  # pow(x, 2) == x*x exactly in IEEE double
    for w, (a, b) in enumerate(zip(start, stop)):
        acc = np.zeros(data.shape[1], dtype=np.float64)
        for r in range(a, b):
            acc += sq[r]
        out[w] = _log(acc / float(b - a) + EPS_LOG)
    return out


def compute_log_power_features_fast(data: np.ndarray, sr: int, window_length: float, window_shift: float) -> np.ndarray:
    """Same function, vectorised over windows (cumulative-sum free: strided gather + sequential-order sum).

    Bit-identical to `compute_log_power_features` when every window has the same length (all configured sizes),
    because the per-window accumulation order (row a, a+1, ...) is kept.  Used for the large parity cases.
    """
    data = np.asarray(data, dtype=np.float64)
    start, stop = window_indices(data.shape[0], sr, window_length, window_shift)
    if len(start) == 0:
        return np.empty((0, data.shape[1]), dtype=np.float64)
    lens = stop - start
    if not np.all(lens == lens[0]):
        return compute_log_power_features(data, sr, window_length, window_shift)
    L = int(lens[0])
    sq = data * data
    acc = np.zeros((len(start), data.shape[1]), dtype=np.float64)
    for k in range(L):
        acc += sq[start + k]
    return _log(acc / float(L) + EPS_LOG)


class WarmStartFrameBuffer:
    """NumPy restatement of the reference's overlap buffer (hga_optimized.pyx:50-131).

    (The compiled reference class aborts under Cython 3.3 when a returned slice is released — SURVEY.md §8c — so the
    streaming oracle uses this restatement; tests/golden/make_golden.py cross-checks it against the compiled class
    in a keep-alive subprocess.)
    """

    def __init__(self, frame_length: float, frame_shift: float, fs: int, nb_channels: int):
        f32 = np.float32
        frame_shift_in_samples = int(f32(frame_shift) * f32(fs))            # :76  <int>(float * int)
        self.frame_length_in_samples = int(f32(frame_length) * f32(fs))     # :77
        self.overlap = self.frame_length_in_samples - frame_shift_in_samples  # :78
        self.first_frame = True                                              # :79
        self.nb_channels = nb_channels
        self.remainder_data = np.zeros((self.overlap, nb_channels), dtype=np.float64)  # :82-88

    def reset(self) -> None:                                                 # :90-94
        self.first_frame = True
        self.remainder_data = np.zeros((self.overlap, self.nb_channels), dtype=np.float64)

    def insert(self, data: np.ndarray) -> np.ndarray:                        # :96-131
        data = np.asarray(data, dtype=np.float64)
        n = data.shape[0]
        if self.first_frame and n >= self.frame_length_in_samples:           # case 1 :104-107
            self.first_frame = False
            self.remainder_data = data[data.shape[0] - self.overlap:, :]
            return data
        if self.first_frame:                                                 # case 2 :111-122
            out = np.zeros((self.frame_length_in_samples, data.shape[1]), dtype=np.float64)
            out[self.frame_length_in_samples - n:, :] = data
            self.first_frame = False
        else:                                                                # case 3 :123-131
            out = np.empty((self.overlap + n, data.shape[1]), dtype=np.float64)
            out[: self.overlap, :] = self.remainder_data
            out[self.overlap:, :] = data
        self.remainder_data = out[out.shape[0] - self.overlap:, :]
        return out


# ----------------------------------------------------------------------------------------------------------------------
# Normalisation and baseline statistics
# ----------------------------------------------------------------------------------------------------------------------
class ZScoreNormalization:
    """(data - means) / stds, broadcast over frames (local/common.py:367-376)."""

    def __init__(self, channel_means: np.ndarray, channel_stds: np.ndarray):
        self.channel_means = channel_means
        self.channel_stds = channel_stds

    def __call__(self, data):
        return (data - self.channel_means) / self.channel_stds


def baseline_statistics(trial_features: Sequence[np.ndarray]) -> Tuple[np.ndarray, np.ndarray]:
    """Per-channel mean and population std over the concatenated frames of a day's baseline trials
    (prepare_corpus.py:190-199; baseline_offline.py:45-55)."""
    cat = np.concatenate(list(trial_features))
    return np.mean(cat, axis=0), np.std(cat, axis=0)


def chan_merge(a: Tuple[np.ndarray, np.ndarray, np.ndarray], b: Tuple[np.ndarray, np.ndarray, np.ndarray]):
    """Chan et al. pairwise merge of (n, mean, M2) — the host/multi-rank combine the build uses."""
    na, ma, m2a = a
    nb, mb, m2b = b
    n = na + nb
    with np.errstate(invalid="ignore", divide="ignore"):
        w = np.where(n > 0, nb / np.where(n > 0, n, 1), 0.0)
    d = mb - ma
    return n, ma + d * w, m2a + m2b + d * d * na * w


# ----------------------------------------------------------------------------------------------------------------------
# The extractor                                                                               local/units.py:97-161
# ----------------------------------------------------------------------------------------------------------------------
class HighGammaExtractor:
    """Behavioural restatement of `local.units.HighGammaExtractor` (local/units.py:97-161) without mne/ezmsg."""

    def __init__(self, fs, nb_electrodes, window_length=0.05, window_shift=0.01, l_freq: int = 70, h_freq: int = 170,
                 pre_transforms: Optional[List[Callable]] = None, post_transforms: Optional[List[Callable]] = None,
                 log_power: Callable = compute_log_power_features_fast):
        self.fs = fs
        self.nb_electrodes = nb_electrodes
        self.window_length = window_length
        self.window_shift = window_shift
        self.model_order = 4                                                   # :108 (unused in the reference too)
        self.step_size = 5                                                     # :109
        self.pre_transform = self._compose_functions(*pre_transforms) if pre_transforms is not None else None
        self.post_transform = self._compose_functions(*post_transforms) if post_transforms is not None else None
        self.framebuffer = WarmStartFrameBuffer(window_length, window_shift, fs, nb_electrodes)   # :112-113
        self.hg_filter, self.fh_filter = design_hga_filters(fs, l_freq, h_freq)                   # :124-126
        self.hg_state = tiled_zi(self.hg_filter, nb_electrodes)                                   # :128-132
        self.fh_state = tiled_zi(self.fh_filter, nb_electrodes)
        self._log_power = log_power

    @staticmethod
    def _compose_functions(*functions):                                        # :134-136
        return reduce(lambda f, g: lambda x: g(f(x)), functions, lambda x: x)

    def filtered(self, data: np.ndarray) -> np.ndarray:
        """The cascade output only (local/units.py:151-152); exposed for kernel-level parity tests."""
        data, self.hg_state = sosfilt(self.hg_filter, data, axis=0, zi=self.hg_state)
        data, self.fh_state = sosfilt(self.fh_filter, data, axis=0, zi=self.fh_state)
        return data

    def extract_features(self, data: np.ndarray) -> np.ndarray:                # :145-161
        if self.pre_transform is not None:
            data = self.pre_transform(data)
        data = self.filtered(data)
        data = self.framebuffer.insert(data)
        data = self._log_power(data, self.fs, self.window_length, self.window_shift)
        if self.post_transform is not None:
            data = self.post_transform(data)
        return data


def extract_offline(x: np.ndarray, fs: int, mu: Optional[np.ndarray] = None, sigma: Optional[np.ndarray] = None,
                    segments: Optional[Iterable[Tuple[int, int]]] = None) -> List[np.ndarray]:
    """One fresh extractor per segment, single call each (prepare_corpus.py:47-50; baseline_offline.py:47-50),
    then the offline z-score applied outside the extractor (prepare_corpus.py:226-227)."""
    x = np.asarray(x, dtype=np.float64)
    segments = [(0, x.shape[0])] if segments is None else list(segments)
    out = []
    for a, b in segments:
        f = HighGammaExtractor(fs, x.shape[1]).extract_features(x[a:b])
        if mu is not None:
            f = (f - mu) / sigma
        out.append(f)
    return out


# ----------------------------------------------------------------------------------------------------------------------
# Seeded synthetic ECoG                                                                          SURVEY.md §8(d)
# ----------------------------------------------------------------------------------------------------------------------
def synthetic_ecog(T: int, C: int, fs: int = 1000, kind: str = "white", seed: int = 20241104, session: int = 0) -> np.ndarray:
    """float64 (T, C) row-major. 'white' = N(0,1); 'ecog' = 1/f noise at 30 uV RMS + 60 Hz (20 uV) + 120 Hz (5 uV)."""
    rng = np.random.Generator(np.random.Philox(key=[seed, session]))
    if kind == "white":
        return rng.standard_normal((T, C))
    if kind == "ecog":
        w = rng.standard_normal((T, C))
        spec = np.fft.rfft(w, axis=0)
        f = np.fft.rfftfreq(T, 1.0 / fs)
        f[0] = f[1] if len(f) > 1 else 1.0
        pink = np.fft.irfft(spec / np.sqrt(f)[:, None], n=T, axis=0)
        pink *= 30.0 / np.sqrt(np.mean(pink ** 2, axis=0, keepdims=True))
        t = np.arange(T)[:, None] / fs
        return np.ascontiguousarray(pink + 20.0 * np.sin(2 * np.pi * 60 * t) + 5.0 * np.sin(2 * np.pi * 120 * t + 0.3))
    raise ValueError(kind)


# ----------------------------------------------------------------------------------------------------------------------
# Channel-axis transforms either side of the path: numpy restatements used by the tests to
# check the channel-map kernel.  Pinned against the reference's own classes, executed from /root/reference/local/common.py
# by tests/golden/make_golden.py (tests/golden/transforms.npz).
# ----------------------------------------------------------------------------------------------------------------------
def select_columns(data: np.ndarray, mapping: Sequence[int]) -> np.ndarray:
    """SelectElectrodesFromBothGrids / SelectElectrodesOverSpeechAreas.__call__ (local/common.py:31-32, 54-55)."""
    return data[:, np.asarray(mapping)]


def car_masks(exclude_channels: Sequence[int], grids: Sequence[np.ndarray], layout: np.ndarray):
    """CommonAverageReferencing.__init__ (local/common.py:313-334): per grid, where to apply and where to average."""
    layout = np.asarray(layout)
    appl = [np.isin(layout, g) for g in grids]
    comp = []
    for g, a in zip(grids, appl):
        c = a.copy()
        for ch in exclude_channels:
            if ch in g:
                c[np.argmax(layout == ch)] = False
        comp.append(c)
    return appl, comp


def common_average_reference(data: np.ndarray, appl: Sequence[np.ndarray], comp: Sequence[np.ndarray]) -> np.ndarray:
    """CommonAverageReferencing.__call__ (local/common.py:336-343): means come from the input, per time sample."""
    out = data.copy()
    for a, c in zip(appl, comp):
        out[:, a] = out[:, a] - np.mean(data[:, c], axis=1)[:, None]
    return out


def bad_channel_patches(bad_channels: Sequence[int], grids: Sequence[np.ndarray], layout: np.ndarray):
    """BadChannelCorrection.__init__ (local/common.py:224-246,263-277): (column of the bad channel, columns of its
    non-bad 8-neighbours on the same grid, in row-major grid order)."""
    layout = np.asarray(layout)
    patches = []
    for ch in bad_channels:
        grid = next(np.asarray(g) for g in grids if ch in g)
        good = ~np.isin(grid, list(bad_channels))
        (r,), (c,) = np.where(grid == ch)
        near = np.zeros(grid.shape, dtype=bool)
        near[max(r - 1, 0):r + 2, max(c - 1, 0):c + 2] = True
        near[r, c] = False
        neigh = grid[near & good]
        patches.append((np.where(layout == ch)[0], np.concatenate([np.where(layout == n)[0] for n in neigh])))
    return patches


def bad_channel_correction(data: np.ndarray, patches) -> np.ndarray:
    """BadChannelCorrection.__call__ (local/common.py:279-284)."""
    out = data.copy()
    for loc, neigh in patches:
        out[:, loc] = np.mean(data[:, neigh], axis=1).reshape((len(data), -1))
    return out


# ----------------------------------------------------------------------------------------------------------------------
# Wire packets at the streaming seam.  Pinned against the reference's own ZMQConnector.interpret_bytes
# and BCI2000Package.pack, executed by tests/golden/make_golden.py (tests/golden/packets.npz).
# ----------------------------------------------------------------------------------------------------------------------
def interpret_bytes(data: bytes) -> np.ndarray:
    """ZMQConnector.interpret_bytes (local/units.py:78-82): '=BBB HH' header, float32 [n_channels][n_samples] payload ->
    (n_samples, n_channels) float64, C order."""
    import struct
    header = struct.Struct('=BBB HH')
    _, _, _, n_channels, n_samples = header.unpack(data[:header.size])
    array = np.frombuffer(data[header.size:], dtype=np.float32).reshape(n_channels, n_samples)
    return np.transpose(array).astype(np.float64, order='C', copy=True)


def pack_packet(payload: np.ndarray) -> bytes:
    """BCI2000Package(nb_channels, nb_samples).pack(payload) (development_amplifier.py:14-25): header (4, 1, 2, C, n) and
    the payload flattened in C order as float32."""
    import struct
    payload = np.asarray(payload)
    return struct.pack('=BBB HH', 4, 1, 2, payload.shape[0], payload.shape[1]) + np.ascontiguousarray(payload, dtype=np.float32).tobytes()


def scaled_signals(signal, gain):
    """`BCI2000MatFile.signals()` (/root/reference/local/common.py:456-459): `signal * gain` with numpy's promotion --
    raw counts (any real dtype) times the per-channel float64 `SourceChGain` vector (or a scalar)."""
    return np.asarray(signal) * np.asarray(gain)

