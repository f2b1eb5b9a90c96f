"""GPU parity: the CUDA path (through the C ABI, via the Python host layer) against the CPU oracle on identical seeded
inputs.  Tolerances (stated here, per feature = per (frame, channel) element):
  * 'exact' precision and every streaming push <= 1024 rows: the arithmetic is the reference's, only the final log()
    may differ by one ulp  ->  max-abs <= 4e-15, max-rel <= 4e-15 on features (|f| < 8).
  * 'f64' precision (fused multiply-adds, hop-block sums, chunked scan with the filter state carried to 1e-11): max-abs
    <= 1e-9 on features, <= 5e-9 on z-scores (division by sigma ~ 0.4); relative error <= 1e-8 where |f| > 1e-2; also on
    input with a DC offset 10^3 times the in-band amplitude (the carry propagates states, it does not re-estimate them).
  * filtered samples (hga_sosfilt) and filter states: bit-identical to scipy.signal.sosfilt.
  * baseline mu / sigma: relative error <= 1e-12.
  * 'f32' precision (float32 cascade arithmetic and power sums): max-abs <= 1e-4 on features (observed ~2e-5),
    <= 5e-4 on z-scores; at 2 kHz (notch poles at radius 0.9988) <= 5e-4 on features.
  * 'mixed' precision (float32 band-pass, float64 notch): max-abs <= 5e-5 on features at 1 kHz, <= 2e-4 at 2 kHz.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import hga_oracle as O  # noqa: E402  (the checker; never on the product path)
from oracle import c_oracle as CO  # noqa: E402

TOL_EXACT = 4e-15
TOL_F64 = 1e-9
TOL_F32 = 1e-4
TOL_MIX = 5e-5


def _oracle_offline(x, fs, mu=None, sigma=None, segments=None):
    """Fresh extractor per segment, one call each, offline z-score — via the C restatement (bit-identical to the
    NumPy oracle, asserted in tests/test_oracle.py)."""
    bp, nt = O.design_hga_filters(fs)
    from scipy.signal import sosfilt_zi
    L, hop = int(0.05 * fs), int(0.01 * fs)
    segments = [(0, len(x))] if segments is None else segments
    return np.concatenate([CO.extract_fused(x[a:b], bp, sosfilt_zi(bp), nt, sosfilt_zi(nt), L, hop, mu=mu, sigma=sigma)
                           for a, b in segments])


def _err(a, b):
    d = np.abs(a - b)
    return d.max(), (d / np.maximum(np.abs(b), 1e-2)).max()


@pytest.fixture(scope="module")
def torch():
    import torch
    return torch


# ------------------------------------------------------------------------------------------------------------ sosfilt
@pytest.mark.parametrize("which", ["bp", "notch"])
def test_sosfilt_bit_exact(which):
    from scipy.signal import sosfilt
    from delayed_speech_synthesis_b200 import ops
    fs, T, C = 1000, 3000, 37
    x = O.synthetic_ecog(T, C, fs, "ecog")
    sos = O.design_hga_filters(fs)[0 if which == "bp" else 1]
    zi = O.tiled_zi(sos, C)
    y_ref, zf_ref = sosfilt(sos, x, axis=0, zi=zi)
    y, zf = ops.sosfilt(sos, x, zi)
    assert np.array_equal(y, y_ref)
    assert np.array_equal(zf, zf_ref)


# ---------------------------------------------------------------------------------------------------------- log power
@pytest.mark.parametrize("rows,C,fs", [(1000, 4, 1000), (50, 1, 1000), (257, 33, 1000), (1000, 8, 2000), (49, 3, 1000)])
def test_log_power_matches_reference_function(rows, C, fs):
    from delayed_speech_synthesis_b200.hga_optimized import compute_log_power_features
    data = O.synthetic_ecog(rows, C, fs, "white", session=3)
    ref = O.compute_log_power_features(data, fs, 0.05, 0.01)
    got = compute_log_power_features(data, fs, 0.05, 0.01)
    assert got.shape == ref.shape and got.dtype == np.float64
    if ref.size:
        assert _err(got, ref)[0] <= TOL_EXACT


@pytest.mark.parametrize("rows,C,fs", [(60_000, 128, 1000), (7_777, 33, 2000), (50, 5, 1000), (1_290, 64, 1000), (200_013, 40, 1000)])
def test_log_power_rolling_sums_are_bit_identical_to_the_window_loop(rows, C, fs):
    """k_frames_roll (one load per sample, five live sums) against k_frames_seq (thread per window; taken for the frame
    buffer's zero-padded geometry) on the same logical rows, and against the oracle's row-order loop: every sum receives
    its rows in the same order, so the bits agree; with the fused z-score too."""
    from delayed_speech_synthesis_b200 import ops
    data = O.synthetic_ecog(rows, C, fs, "ecog", session=6)
    hop = int(0.01 * fs)
    padded = np.concatenate([np.zeros((hop, C)), data])            # rows the zero-padded call sees implicitly
    mu, sigma = np.linspace(1.0, 2.0, C), np.linspace(0.5, 1.5, C)
    roll = ops.log_power(padded, fs, 0.05, 0.01)
    roll_z = ops.log_power(padded, fs, 0.05, 0.01, mu, sigma)
    seq = ops.log_power(data, fs, 0.05, 0.01, zero_rows=hop)
    seq_z = ops.log_power(data, fs, 0.05, 0.01, mu, sigma, zero_rows=hop)
    assert roll.shape == seq.shape == ((rows + hop - int(0.05 * fs)) // hop + 1, C)
    assert np.array_equal(roll, seq) and np.array_equal(roll_z, seq_z)
    n = min(len(roll), 300)
    ref = O.compute_log_power_features(padded[:(n - 1) * hop + int(0.05 * fs)], fs, 0.05, 0.01)
    assert _err(roll[:n], ref)[0] <= TOL_EXACT


def test_log_power_strided_and_errors():
    from delayed_speech_synthesis_b200.hga_optimized import compute_log_power_features
    big = O.synthetic_ecog(400, 12, 1000, "white")
    view = big[::2, 1:9]                                    # non-contiguous: "any strides accepted"
    ref = O.compute_log_power_features(np.ascontiguousarray(view), 1000, 0.05, 0.01)
    assert _err(compute_log_power_features(view, 1000, 0.05, 0.01), ref)[0] <= TOL_EXACT
    with pytest.raises(ValueError):
        compute_log_power_features(big.astype(np.float32), 1000, 0.05, 0.01)
    with pytest.raises(ValueError):
        compute_log_power_features(big[0], 1000, 0.05, 0.01)


def test_log_power_irregular_windows():
    """window_shift = 0.0125 s at 1 kHz: the reference's float32 index arithmetic rounds starts off any regular grid."""
    from delayed_speech_synthesis_b200.hga_optimized import compute_log_power_features
    data = O.synthetic_ecog(700, 5, 1000, "white", session=9)
    ref = O.compute_log_power_features(data, 1000, 0.05, 0.0125)
    got = compute_log_power_features(data, 1000, 0.05, 0.0125)
    assert got.shape == ref.shape
    assert _err(got, ref)[0] <= TOL_EXACT


def test_known_answers():
    """Constant input -> band-pass output decays to 0 -> feature -> log(0.01); a unit step with the unscaled
    sosfilt_zi state is the notch filter's exact steady state."""
    from delayed_speech_synthesis_b200.units import HighGammaExtractor
    ex = HighGammaExtractor(1000, 4, precision="exact")
    f = ex.extract_features(np.full((20000, 4), 3.0))
    assert abs(f[-1] - np.log(0.01)).max() < 1e-9


# ------------------------------------------------------------------------------------------------------------ offline
@pytest.mark.parametrize("precision,tol", [("exact", TOL_EXACT), ("f64", TOL_F64)])
@pytest.mark.parametrize("kind", ["white", "ecog"])
def test_offline_k1_shape(precision, tol, kind):
    """BASELINE config 1: 128 ch x 60 s @ 1 kHz, single call on a fresh extractor."""
    from delayed_speech_synthesis_b200.units import HighGammaExtractor
    fs, T, C = 1000, 60000, 128
    x = O.synthetic_ecog(T, C, fs, kind)
    ref = _oracle_offline(x, fs)
    got = HighGammaExtractor(fs, C, precision=precision).extract_features(x)
    assert got.shape == ref.shape == (5996, C)
    a, r = _err(got, ref)
    assert a <= tol and r <= 10 * tol, (a, r)


@pytest.mark.parametrize("precision,tol", [("exact", TOL_EXACT), ("f64", TOL_F64)])
def test_offline_segments_and_zscore(precision, tol):
    """Trials as in prepare_corpus.py:42-52 (fresh extractor each, ragged lengths incl. the minimum L) + z-score."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C = 1000, 64
    x = O.synthetic_ecog(30000, C, fs, "ecog", session=1)
    segs = [(0, 2540), (2540, 2590), (3000, 9017), (9017, 30000)]
    base = _oracle_offline(x, fs, segments=segs)
    mu, sigma = base.mean(0), base.std(0)
    ref = _oracle_offline(x, fs, mu, sigma, segments=segs)
    eng = HgaEngine(fs, C, precision=precision)
    got = eng.extract_offline(x, segs, mu, sigma)
    assert got.shape == ref.shape
    a, r = _err(got, ref)
    assert a <= 5 * tol, (a, r)


def test_offline_device_tensor_matches_host(torch):
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C = 1000, 128
    x = O.synthetic_ecog(120000, C, fs, "white", session=2)
    eng = HgaEngine(fs, C, precision="f64")
    host = eng.extract_offline(x)
    dev = eng.extract_offline(torch.from_numpy(x).cuda()).cpu().numpy()
    assert np.array_equal(host, dev)
    assert _err(host, _oracle_offline(x, fs))[0] <= TOL_F64


def test_offline_time_chunked_long(torch):
    """10 min x 128 ch: many time-chunks with warm-up; oracle via the multi-threaded C restatement."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, C = 1000, 600000, 128
    x = O.synthetic_ecog(T, C, fs, "white", session=5)
    x += 5.0 * np.sin(2 * np.pi * 120 * np.arange(T) / fs)[:, None]   # line harmonic sitting in the notch
    ref = _oracle_offline(x, fs)
    got = HgaEngine(fs, C, precision="f64").extract_offline(torch.from_numpy(x).cuda()).cpu().numpy()
    a, r = _err(got, ref)
    assert a <= TOL_F64, (a, r)


@pytest.mark.parametrize("cfg,precision", [("k2", "f64"), ("k2", "f32"), ("k2", "mixed"), ("k4", "f64"), ("k4", "f32")])
def test_offline_full_size_windows(cfg, precision, torch):
    """BASELINE configs 2 and 4 at full size (128 ch x 1 h @ 1 kHz = 3.69 GB; 1024 ch x 10 min @ 2 kHz = 9.83 GB), generated
    on the device: the start of the recording against a fresh-start oracle, and windows at seeded positions anywhere in
    it against an oracle run that starts `lead` rows earlier (its own start transient has decayed by 1e-14 by then).
    Checks the time-chunking (warm-up, chunk seams, two-wave plans) at the sizes the bench runs."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, C, lead, span, n_win = {"k2": (1000, 3_600_000, 128, 14000, 1000, 6), "k4": (2000, 1_200_000, 1024, 28000, 1000, 3)}[cfg]
    L, hop = int(0.05 * fs), int(0.01 * fs)
    tol = TOL_F64 if precision == "f64" else (TOL_MIX if precision == "mixed" else (TOL_F32 if fs == 1000 else 5 * TOL_F32))
    g = torch.Generator(device="cuda").manual_seed(20241104)
    x = torch.randn((T, C), dtype=torch.float64, device="cuda", generator=g)
    eng = HgaEngine(fs, C, precision=precision)
    feats = eng.extract_offline(x)
    assert feats.shape == ((T - L) // hop + 1, C)
    plan = eng.last_plan()
    assert plan["carried"] >= 8, plan
    n0 = 60000 if cfg == "k2" else 20000
    ref = _oracle_offline(x[:n0].cpu().numpy(), fs)
    assert _err(feats[:len(ref)].cpu().numpy(), ref)[0] <= tol
    rng = np.random.Generator(np.random.Philox(key=[20241104, 99]))
    Lc = plan["chunk_rows"]
    n_ch = -(-T // Lc)                                                                 # chunks of the one segment
    seams = [k * Lc - span // 2 for k in (-(-(lead + span) // Lc), n_ch // 2, n_ch - 1)]   # windows ON chunk seams
    for t0 in sorted(int(v) * hop for v in rng.integers(lead // hop, (T - span) // hop, size=n_win)) + seams + [T - span]:
        ref = _oracle_offline(x[t0 - lead:t0 + span].cpu().numpy(), fs)      # frame k of the slice = frame (t0 - lead) / hop + k
        k0 = lead // hop
        got = feats[t0 // hop:t0 // hop + (span - L) // hop + 1].cpu().numpy()
        assert _err(got, ref[k0:k0 + len(got)])[0] <= tol, t0
    assert bool(torch.isfinite(feats).all())


@pytest.mark.parametrize("precision,tol,pad", [("f64", TOL_F64, 0), ("f64", TOL_F64, 1), ("f32", TOL_F32, 0), ("mixed", TOL_MIX, 0)])
def test_offline_ragged_segments_every_kernel(precision, tol, pad, torch):
    """The lane kernels (pad = 0) and the generic thread-sequential kernel (pad = 1: an odd leading dimension cannot be a
    TMA tensor) against the oracle: ragged segments (partial tiles, lengths that are not hop multiples, the minimum length L),
    C not a multiple of 32 or 64, z-score."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C = 1000, 70
    x = O.synthetic_ecog(90007, C, fs, "ecog", session=11)
    segs = [(0, 50), (50, 2590), (2600, 2655), (3000, 9017), (9018, 90007)]
    base = _oracle_offline(x, fs, segments=segs)
    mu, sigma = base.mean(0), base.std(0)
    ref = _oracle_offline(x, fs, mu, sigma, segments=segs)
    wide = torch.zeros((len(x), C + pad), dtype=torch.float64, device="cuda")
    wide[:, :C] = torch.from_numpy(x).cuda()
    eng = HgaEngine(fs, C, precision=precision)
    got = eng.extract_offline(wide[:, :C], segs, mu, sigma).cpu().numpy()
    assert got.shape == ref.shape
    assert _err(got, ref)[0] <= 5 * tol
    assert (eng.last_plan()["chunk_rows"] > 0) == (pad == 0)      # which kernel ran


def test_offline_non_standard_cascade_takes_the_generic_kernel(torch):
    """An order-4 band-pass (4 + 8 sections) is outside what the lane kernels are built for: the generic kernel runs the
    caller's sections as given (float64 only; the packed / mixed arithmetic refuses)."""
    from scipy.signal import iirfilter, sosfilt_zi
    from delayed_speech_synthesis_b200 import HgaError
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C = 1000, 40
    x = O.synthetic_ecog(60000, C, fs, "ecog", session=13)
    bp = iirfilter(4, np.array([70, 170]) / (fs / 2), btype="bandpass", ftype="butter", output="sos")
    nt = O.design_hga_filters(fs)[1]
    ref = CO.extract_fused(x, bp, sosfilt_zi(bp), nt, sosfilt_zi(nt), 50, 10)
    eng = HgaEngine(fs, C, precision="f64", sos_bp=bp, sos_notch=nt)
    got = eng.extract_offline(torch.from_numpy(x).cuda()).cpu().numpy()
    assert _err(got, ref)[0] <= TOL_F64
    assert eng.last_plan()["chunk_rows"] == 0
    with pytest.raises(HgaError):
        HgaEngine(fs, C, precision="f32", sos_bp=bp, sos_notch=nt).extract_offline(torch.from_numpy(x).cuda())


@pytest.mark.parametrize("precision,tol", [("f64", TOL_F64), ("f32", 3 * TOL_F32), ("mixed", 5 * TOL_MIX)])
def test_offline_dc_offset_across_chunk_seams(precision, tol, torch):
    """Raw amplifier signals sit on per-channel offsets far above the high-gamma amplitude: +-5000 here against
    an in-band RMS of ~5.  The filter states at a chunk seam scale with the offset; the carry propagates them exactly, so the
    float64 bound holds unchanged.  The float32 band-pass sections (float32 and mixed modes) remove each channel's offset in
    float64 before they convert: tolerance x 3 / x 5, the start transient of the +-5000 step included."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, C = 1000, 400000, 64
    x = O.synthetic_ecog(T, C, fs, "ecog", session=21)
    x += np.linspace(-5000.0, 5000.0, C)[None, :]
    ref = _oracle_offline(x, fs)
    eng = HgaEngine(fs, C, precision=precision)
    got = eng.extract_offline(torch.from_numpy(x).cuda()).cpu().numpy()
    plan = eng.last_plan()
    assert plan["carried"] >= 100, plan                                # the recording really was cut
    a, r = _err(got, ref)
    assert a <= tol, (a, r, plan)
    # the frames around every seam, separately (where a state error would sit)
    Lc, hop = plan["chunk_rows"], 10
    seam = np.concatenate([np.arange(k * Lc // hop - 5, k * Lc // hop + 40) for k in range(1, plan["chunks"])])
    seam = seam[seam < len(ref)]
    assert np.abs(got[seam] - ref[seam]).max() <= tol


def test_offline_f32_weak_signal_on_a_large_offset(torch):
    """float32 arithmetic with an in-band amplitude four to five decades below the channel's offset (sigma 0.1 .. 10 on
    offsets up to +-8000, where one float32 quantum of the raw sample is 4.9e-4): the section-pipelined kernel removes the
    median of each segment's first three samples in float64 before it converts (LaneF32::kOffset), so the plain float32
    bound holds on carried chunks, at the seams and on segments that do not start at row 0.  The first 8 s of a segment are
    the reference's own start transient (zi is not scaled by x[0]: a step of the offset's size rings through the notch);
    float32 follows it to 1e-3 while it decays to the weak signal."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, C = 1000, 300000, 64
    L, hop = 50, 10
    rng = np.random.default_rng(5)
    x = rng.standard_normal((T, C)) * np.logspace(-1, 1, C)[None, :]
    x += np.linspace(-8000.0, 8000.0, C)[None, :]
    x[1, 7] += 40.0                                                     # one stray sample among the three the estimate reads
    eng = HgaEngine(fs, C, precision="f32")
    for segs in ([(0, T)], [(1000, 130000), (130500, 299000), (299000, 299200)]):
        ref = _oracle_offline(x, fs, segments=segs)
        got = eng.extract_offline(torch.from_numpy(x).cuda(), segs).cpu().numpy()
        plan = eng.last_plan()
        assert plan["carried"] >= 50, plan
        start = np.zeros(len(ref), dtype=bool)                          # frames inside a segment's first 8 s
        at = 0
        for a, b in segs:
            n = (b - a - L) // hop + 1
            start[at:at + min(n, 800)] = True
            at += n
        assert at == len(ref)
        d = np.abs(got - ref)
        assert d[~start].max() <= TOL_F32, (d[~start].max(), plan, segs)
        assert d[start].max() <= 10 * TOL_F32, (d[start].max(), plan, segs)
    # the host-buffer call runs the same items slab by slab
    d = np.abs(eng.extract_offline(x) - _oracle_offline(x, fs))
    assert d[800:].max() <= TOL_F32 and d[:800].max() <= 10 * TOL_F32


@pytest.mark.parametrize("C", [6, 130])
def test_offline_sequential_carry_ragged_channels_two_segments(C, torch):
    """One-minute segments are cut into chunks of a few hundred rows, far shorter than the carry: the start states come from
    the sequential chain (`k_lane_carry`: CTAs of 4 adjacent channels, cp.async ring of end-state tiles), here with a channel
    count that is not a multiple of 4 or of a warp's channel group, and two segments of different chunk counts."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T = 1000, 100_000
    x = O.synthetic_ecog(T, C, fs, "ecog", session=71) + np.linspace(-300.0, 300.0, C)[None, :]
    segs = [(0, 61_003), (61_500, T)]
    ref = _oracle_offline(x, fs, segments=segs)
    for precision, tol in (("f64", TOL_F64), ("f32", TOL_F32)):
        eng = HgaEngine(fs, C, precision=precision)
        got = eng.extract_offline(torch.from_numpy(x).cuda(), segs).cpu().numpy()
        plan = eng.last_plan()
        assert plan["carried"] >= 40 and plan["chunk_rows"] < 2500, plan       # window far beyond 8 chunks: the sequential chain
        assert got.shape == ref.shape and _err(got, ref)[0] <= tol, (precision, _err(got, ref), plan)


def test_offline_exact_overlapping_segments(torch):
    """Trials closer than the 40 ms tail overlap in x (prepare_corpus.py:49); every one is filtered from its own fresh
    state, in every precision (the exact path used to share one scratch row per x row)."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C = 1000, 33
    x = O.synthetic_ecog(9000, C, fs, "ecog", session=17)
    segs = [(0, 2540), (2500, 5040), (2520, 2600), (5000, 9000), (100, 8000)]
    ref = _oracle_offline(x, fs, segments=segs)
    for precision, tol in (("exact", TOL_EXACT), ("f64", TOL_F64), ("f32", TOL_F32)):
        got = HgaEngine(fs, C, precision=precision).extract_offline(torch.from_numpy(x).cuda(), segs).cpu().numpy()
        assert got.shape == ref.shape and _err(got, ref)[0] <= tol, precision
        assert _err(HgaEngine(fs, C, precision=precision).extract_offline(x, segs), ref)[0] <= tol, precision


@pytest.mark.parametrize("fs,kind,tol", [(1000, "white", TOL_MIX), (1000, "ecog", TOL_MIX), (2000, "white", 4 * TOL_MIX)])
def test_offline_mixed_mode(fs, kind, tol, torch):
    """float32 band-pass + float64 notch: stated tolerance 5e-5 per feature at 1 kHz, 2e-4 at 2 kHz."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    T, C = 200 * fs, 128
    x = O.synthetic_ecog(T, C, fs, kind, session=14)
    ref = _oracle_offline(x, fs)
    mu, sigma = ref.mean(0), ref.std(0)
    eng = HgaEngine(fs, C, precision="mixed")
    got = eng.extract_offline(torch.from_numpy(x).cuda()).cpu().numpy()
    a, r = _err(got, ref)
    assert a <= tol, (a, r)
    z = eng.extract_offline(torch.from_numpy(x).cuda(), None, mu, sigma).cpu().numpy()
    assert np.abs(z - (ref - mu) / sigma).max() <= 5 * tol


def test_offline_pageable_and_pinned_host_buffers_agree(torch):
    """hga_extract_offline_host reads page-locked memory where it lies and stages pageable memory (what numpy hands over)
    through its own pinned buffers: same bits either way, strided input and output included."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, C = 1000, 300000, 128
    x = O.synthetic_ecog(T, C + 2, fs, "white", session=19)[:, 1:C + 1]        # pageable, leading dimension C + 2
    eng = HgaEngine(fs, C, precision="f64")
    a = eng.extract_offline(x)
    xp = torch.from_numpy(np.ascontiguousarray(x)).pin_memory()
    outp = torch.empty((eng.num_frames(T), C), dtype=torch.float64).pin_memory()
    b = eng.extract_offline(xp.numpy(), out=outp.numpy())
    assert np.array_equal(a, b)
    assert _err(a, _oracle_offline(np.ascontiguousarray(x), fs))[0] <= TOL_F64


@pytest.mark.parametrize("kind", ["white", "ecog"])
def test_offline_f32_mode(kind, torch):
    """float32 cascade arithmetic: stated tolerance 1e-4 max-abs per feature (log-power), 5e-4 per z-score."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, C = 1000, 200000, 128
    x = O.synthetic_ecog(T, C, fs, kind, session=12)
    ref = _oracle_offline(x, fs)
    mu, sigma = ref.mean(0), ref.std(0)
    eng = HgaEngine(fs, C, precision="f32")
    got = eng.extract_offline(torch.from_numpy(x).cuda()).cpu().numpy()
    a, r = _err(got, ref)
    assert a <= TOL_F32, (a, r)
    z = eng.extract_offline(torch.from_numpy(x).cuda(), None, mu, sigma).cpu().numpy()
    assert np.abs(z - (ref - mu) / sigma).max() <= 5 * TOL_F32
    assert np.array_equal(eng.extract_offline(x), got)          # host-buffer path == device path


def test_offline_2khz_high_density_prefix(torch):
    """BASELINE config 4 geometry (2 kHz: L = 100, hop = 20) on a 1024-channel, 30 s prefix."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, C = 2000, 60000, 1024
    x = O.synthetic_ecog(T, C, fs, "white", session=6)
    ref = _oracle_offline(x, fs)
    got = HgaEngine(fs, C, precision="f64").extract_offline(torch.from_numpy(x).cuda()).cpu().numpy()
    assert got.shape == ref.shape == (2996, C)
    assert _err(got, ref)[0] <= TOL_F64
    # float32 cascade at 2 kHz: the notch poles sit at radius 0.9988 (0.9976 at 1 kHz), rounding noise is amplified more;
    # stated tolerance at this rate 5e-4 max-abs per feature (observed 2.6e-4 on white noise)
    got32 = HgaEngine(fs, C, precision="f32").extract_offline(torch.from_numpy(x).cuda()).cpu().numpy()
    assert _err(got32, ref)[0] <= 5 * TOL_F32


def test_offline_rejects_short_segment():
    from delayed_speech_synthesis_b200 import HgaError
    from delayed_speech_synthesis_b200.engine import HgaEngine
    eng = HgaEngine(1000, 8)
    with pytest.raises(HgaError):
        eng.extract_offline(np.zeros((49, 8)))


# ---------------------------------------------------------------------------------------------------------- streaming
@pytest.mark.parametrize("chunks", [[10] * 60, [40] * 20, [20, 10, 10, 40, 70, 10, 30], [60, 10, 10, 25, 13, 50], [7, 9, 33, 10, 10]])
def test_streaming_matches_oracle_unit(chunks):
    """Long-lived extractor fed packet by packet (decode_online.py): all three frame-buffer cases, ragged packets that
    drift off the hop grid, fused z-score.  Arithmetic is reference-order -> ulp-level agreement."""
    from delayed_speech_synthesis_b200.common import ZScoreNormalization
    from delayed_speech_synthesis_b200.units import HighGammaExtractor
    fs, C = 1000, 64
    x = O.synthetic_ecog(sum(chunks), C, fs, "ecog", session=4)
    rng = np.random.default_rng(0)
    mu, sigma = rng.normal(3, 1, (1, C)), rng.uniform(0.3, 0.6, (1, C))
    ref_ex = O.HighGammaExtractor(fs, C, post_transforms=[O.ZScoreNormalization(mu, sigma)], log_power=O.compute_log_power_features)
    ex = HighGammaExtractor(fs, C, post_transforms=[ZScoreNormalization(mu, sigma)])
    t = 0
    for n in chunks:
        ref = ref_ex.extract_features(x[t:t + n])
        got = ex.extract_features(x[t:t + n])
        assert got.shape == ref.shape, (n, got.shape, ref.shape)
        if ref.size:
            assert _err(got, ref)[0] <= 10 * TOL_EXACT
        t += n
    assert np.array_equal(ex.hg_state, ref_ex.hg_state)
    assert np.array_equal(ex.fh_state, ref_ex.fh_state)


def test_streaming_device_and_host_pushes_interleave(torch):
    """Device pushes run on the caller's stream, host pushes on the handle's own: the library orders one behind the other
    (no synchronise in between here), so a mixed sequence equals the oracle's packet-by-packet result."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C, n, P = 1000, 64, 10, 120
    x = O.synthetic_ecog(n * P, C, fs, "ecog", session=9)
    ref_ex = O.HighGammaExtractor(fs, C, log_power=O.compute_log_power_features)
    eng = HgaEngine(fs, C, precision="exact")
    xd = torch.from_numpy(x).cuda()
    side = torch.cuda.Stream()
    got, want = [], []
    for i in range(P):
        want.append(ref_ex.extract_features(x[i * n:(i + 1) * n]))
        if i % 3 == 1:
            with torch.cuda.stream(side):                    # a non-default caller stream, never synchronised by the test
                got.append(eng.push(xd[i * n:(i + 1) * n]))
        else:
            got.append(eng.push(x[i * n:(i + 1) * n]))
    torch.cuda.synchronize()
    for g, w in zip(got, want):
        g = g.cpu().numpy() if hasattr(g, "cpu") else g
        assert g.shape == w.shape and _err(g, w)[0] <= 10 * TOL_EXACT


def test_streaming_long_lived_handle_10500_packets():
    """BASELINE config 3 / SURVEY 8d: >= 10 500 ten-sample pushes on ONE handle (105 s of signal).  Frames within the
    reference-order bound of the oracle's filtering + framing of the whole signal (streaming == offline on
    [40 zero rows; y], SURVEY 8c pin (i)); the carried filter states bit-equal to scipy's after the last packet."""
    from scipy.signal import sosfilt, sosfilt_zi
    from delayed_speech_synthesis_b200.common import ZScoreNormalization
    from delayed_speech_synthesis_b200.units import HighGammaExtractor
    fs, C, n_pk, n = 1000, 128, 10500, 10
    x = O.synthetic_ecog(n_pk * n, C, fs, "ecog", session=23)
    rng = np.random.default_rng(1)
    mu, sigma = rng.normal(3, 1, (1, C)), rng.uniform(0.3, 0.6, (1, C))
    ex = HighGammaExtractor(fs, C, post_transforms=[ZScoreNormalization(mu, sigma)])
    frames = np.concatenate([ex.extract_features(x[i * n:(i + 1) * n]) for i in range(n_pk)])
    bp, nt = O.design_hga_filters(fs)
    y, z_bp = sosfilt(bp, x, axis=0, zi=O.tiled_zi(bp, C))
    y, z_nt = sosfilt(nt, y, axis=0, zi=O.tiled_zi(nt, C))
    ref = (O.compute_log_power_features_fast(np.concatenate([np.zeros((40, C)), y]), fs, 0.05, 0.01) - mu) / sigma
    assert frames.shape == ref.shape == (n_pk, C)
    assert _err(frames, ref)[0] <= 25 * TOL_EXACT          # z-score divides the ulp of log() by sigma >= 0.3
    assert np.array_equal(ex.hg_state, z_bp) and np.array_equal(ex.fh_state, z_nt)


def test_streaming_equals_offline_with_zero_pad():
    """SURVEY §8c pin (i): 10-sample pushes == offline framing of [40 zero rows ; filtered signal]."""
    from delayed_speech_synthesis_b200.units import HighGammaExtractor
    from delayed_speech_synthesis_b200 import ops
    fs, C, n = 1000, 16, 10
    x = O.synthetic_ecog(3000, C, fs, "white", session=8)
    ex = HighGammaExtractor(fs, C)
    stream = np.concatenate([ex.extract_features(x[t:t + n]) for t in range(0, len(x), n)])
    oe = O.HighGammaExtractor(fs, C)
    y = oe.filtered(x)
    off = ops.log_power(y, fs, 0.05, 0.01, zero_rows=40)
    assert stream.shape == off.shape == (300, C)
    assert np.array_equal(stream, off)


@pytest.mark.parametrize("precision,tol", [("exact", TOL_EXACT), ("f64", TOL_F64)])
def test_bulk_push_carries_state(precision, tol, torch):
    """Large pushes (bulk path) interleaved with packets: state, remainder and frame grid carry across both paths."""
    from delayed_speech_synthesis_b200.units import HighGammaExtractor
    fs, C = 1000, 32
    chunks = [5000, 10, 10, 30000, 40, 2003, 10]
    x = O.synthetic_ecog(sum(chunks), C, fs, "ecog", session=7)
    ref_ex, ex = O.HighGammaExtractor(fs, C), HighGammaExtractor(fs, C, precision=precision)
    t = 0
    for n in chunks:
        ref, got = ref_ex.extract_features(x[t:t + n]), ex.extract_features(x[t:t + n])
        assert got.shape == ref.shape, (n, got.shape, ref.shape)
        assert _err(got, ref)[0] <= tol, n
        t += n
    np.testing.assert_allclose(ex.fh_state, ref_ex.fh_state, rtol=0, atol=1e-9 if precision == "f64" else 0)


def test_state_roundtrip_and_reset():
    from delayed_speech_synthesis_b200.units import HighGammaExtractor
    fs, C = 1000, 8
    x = O.synthetic_ecog(200, C, fs, "white")
    a, b = HighGammaExtractor(fs, C), HighGammaExtractor(fs, C)
    for t in range(0, 100, 10):
        a.extract_features(x[t:t + 10])
    hg, fh, rem, ff = a._engine.get_state()
    b._engine.set_state(hg, fh, rem, ff)
    assert np.array_equal(a.extract_features(x[100:110]), b.extract_features(x[100:110]))
    assert np.array_equal(HighGammaExtractor(fs, C).hg_state, O.tiled_zi(a.hg_filter, C))
    a._engine.reset()
    fresh = HighGammaExtractor(fs, C)
    assert np.array_equal(a.extract_features(x[:10]), fresh.extract_features(x[:10]))


def test_unit_message_interface():
    import asyncio
    from delayed_speech_synthesis_b200.ezmsg_compat import TimeSeriesMessage
    from delayed_speech_synthesis_b200.units import HighGammaActivity, HighGammaActivitySettings
    unit = HighGammaActivity(HighGammaActivitySettings(fs=1000, nb_electrodes=16))
    unit.setup()
    x = O.synthetic_ecog(40, 16, 1000, "white")

    async def run():
        return [m async for m in unit.process(TimeSeriesMessage(data=x, fs=1000.0))]
    (stream, msg), = asyncio.run(run())
    assert stream is unit.OUTPUT and msg.fs == 100.0 and msg.data.shape == (1, 16)
    ref = O.HighGammaExtractor(1000, 16).extract_features(x)
    assert _err(msg.data, ref)[0] <= 10 * TOL_EXACT


# --------------------------------------------------------------------------------------------------------- statistics
def test_baseline_statistics_and_merge(torch):
    from delayed_speech_synthesis_b200.ops import ChannelStats
    rng = np.random.default_rng(1)
    trials = [rng.normal(3.0, 0.4, (n, 128)) + np.linspace(-2, 6, 128) for n in (250, 311, 1, 4096, 77)]
    mu_ref, sd_ref = O.baseline_statistics(trials)
    st = ChannelStats(128)
    for t in trials:
        st.update(t)
    mu, sd = st.finalize()
    np.testing.assert_allclose(mu, mu_ref, rtol=1e-12, atol=0)
    np.testing.assert_allclose(sd, sd_ref, rtol=1e-12, atol=0)
    # two "ranks": merged accumulators == single accumulator
    a, b = ChannelStats(128), ChannelStats(128)
    for t in trials[:2]:
        a.update(t)
    for t in trials[2:]:
        b.update(t)
    mu2, sd2 = a.merge(b).finalize()
    np.testing.assert_allclose(mu2, mu_ref, rtol=1e-12, atol=0)
    np.testing.assert_allclose(sd2, sd_ref, rtol=1e-12, atol=0)


def test_zscore_callable_matches_numpy():
    from delayed_speech_synthesis_b200.common import ZScoreNormalization
    rng = np.random.default_rng(2)
    f, mu, sd = rng.normal(size=(500, 64)), rng.normal(size=64), rng.uniform(0.2, 2, 64)
    assert np.array_equal(ZScoreNormalization(mu, sd)(f), (f - mu) / sd)


@pytest.mark.parametrize("F,C", [(4099, 128), (517, 100), (33, 1024), (300, 6)])
def test_zscore_quad_and_scalar_kernels_are_bit_exact(F, C):
    """k_zscore4 (C % 4 == 0: four channels per thread, 256-bit loads / stores) and k_zscore (any C) against numpy's
    (f - mu) / sigma, bit for bit (/root/reference/local/common.py:375-376); a row-padded input too."""
    import torch
    from delayed_speech_synthesis_b200 import ops
    rng = np.random.default_rng(5)
    f, mu, sd = rng.normal(size=(F, C)), rng.normal(size=C), rng.uniform(0.2, 2, C)
    fd, mud, sdd = (torch.from_numpy(a).cuda() for a in (f, mu, sd))
    assert np.array_equal(ops.zscore(fd, mud, sdd).cpu().numpy(), (f - mu) / sd)
    wide = torch.zeros((F, C + 4), dtype=torch.float64, device="cuda")
    wide[:, :C] = fd
    assert np.array_equal(ops.zscore(wide[:, :C], mud, sdd).cpu().numpy(), (f - mu) / sd)
    odd = torch.zeros((F, C + 3), dtype=torch.float64, device="cuda")
    odd[:, :C] = fd
    assert np.array_equal(ops.zscore(odd[:, :C], mud, sdd).cpu().numpy(), (f - mu) / sd)
