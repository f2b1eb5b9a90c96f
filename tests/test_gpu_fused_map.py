"""The pre-transform chain inside the extraction kernel's load stage (`hga_extract_offline_mapped`):
`HighGammaExtractor.extract_features` = pre_transform -> filters -> frames (/root/reference/local/units.py:147-161) with the
chains of prepare_corpus.py:152-160 (select both grids -> per-grid common average -> select the speech grid), against the
oracle's transforms + extraction on the same input.  Tolerances: packed float32 arithmetic 1e-4 per feature (5e-4 on z-scores),
as in test_gpu_parity.py; the fused and the two-pass float32 results agree to 2e-5 (they differ only in the summation order of
the group means, before the rounding to float32)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import hga_oracle as O  # noqa: E402

from test_gpu_parity import TOL_F32, TOL_F64, _oracle_offline  # noqa: E402


def _chain(c_raw, seed):
    """Raw columns -> 128 grid channels (a permutation that drops the rest) -> CAR per 8x8 grid with excluded channels ->
    the 64 channels of the second grid in another order.  Returns (oracle callables, the map spec's ingredients)."""
    from delayed_speech_synthesis_b200 import transforms as TR
    rng = np.random.default_rng(seed)
    both = rng.permutation(c_raw)[:128]
    layout = np.arange(128) + 1
    grids = [np.flip(layout[:64].reshape(8, 8), axis=0), layout[64:].reshape(8, 8)]
    bad = [3, 19, 38, 77, 120]
    appl, comp = O.car_masks(bad, grids, layout)
    speech = 64 + rng.permutation(64)
    pre_o = [lambda d: O.select_columns(d, both), lambda d: O.common_average_reference(d, appl, comp), lambda d: O.select_columns(d, speech)]
    pre = [TR.ChannelSelection(both), TR.CommonAverageReferencing(bad, grids, layout), TR.ChannelSelection(speech)]
    return pre_o, pre


@pytest.mark.parametrize("c_raw", [129, 128, 136])
def test_fused_chain_matches_oracle_and_two_pass(c_raw):
    import torch
    from delayed_speech_synthesis_b200 import _lib, transforms as TR
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T = 1000, 120_037
    raw = O.synthetic_ecog(T, c_raw, fs, "ecog", session=51) + np.linspace(-800.0, 900.0, c_raw)[None, :]   # per-channel offsets: CAR matters
    pre_o, pre = _chain(c_raw, 3)
    mapped = raw
    for f in pre_o:
        mapped = f(mapped)
    segs = [(0, 50_000), (50_010, 50_060), (50_100, 119_999)]
    base = _oracle_offline(mapped, fs, segments=segs)
    mu, sigma = base.mean(0), base.std(0)
    want = _oracle_offline(mapped, fs, mu, sigma, segs)
    spec = TR.fold(pre, c_raw)
    assert spec is not None and spec.c_out == 64
    lib = _lib.load()
    lib.hga_debug_guards(1)
    try:
        cmap = TR.ChannelMap(spec, 0)
        eng = HgaEngine(fs, 64, precision="f32", device=0)
        xd = torch.from_numpy(raw).cuda()
        got = eng.extract_offline_mapped(xd, cmap, segs, mu, sigma)
        torch.cuda.synchronize()
        assert eng.last_fused
        assert lib.hga_debug_check_guards() == 0
        two = eng.extract_offline(cmap.apply(xd), segs, mu, sigma)
        g, t = got.cpu().numpy(), two.cpu().numpy()
    finally:
        lib.hga_debug_guards(0)
    assert g.shape == want.shape and np.isfinite(g).all()
    assert np.abs(g - want).max() <= 5 * TOL_F32
    assert np.abs(g - t).max() <= 2e-5
    assert eng.last_plan()["carried"] > 0                      # several chunks per segment: seams covered


def test_fused_gather_only_two_channel_groups():
    """A pure selection (no averaging group) onto 128 output channels: two items per chunk."""
    import torch
    from delayed_speech_synthesis_b200 import transforms as TR
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, c_raw = 1000, 40_013, 140
    raw = O.synthetic_ecog(T, c_raw, fs, "white", session=52)
    idx = np.random.default_rng(9).permutation(c_raw)[:128]
    want = _oracle_offline(raw[:, idx], fs)
    spec = TR.fold([TR.ChannelSelection(idx)], c_raw)
    eng = HgaEngine(fs, 128, precision="f32", device=0)
    got = eng.extract_offline_mapped(torch.from_numpy(raw).cuda(), TR.ChannelMap(spec, 0)).cpu().numpy()
    assert eng.last_fused and np.abs(got - want).max() <= TOL_F32


def test_mapped_call_without_fusion_is_the_two_pass_result():
    """float64 handles (and anything else the fused kernel does not cover) run the map as its own pass: same entry point,
    float64 tolerance."""
    import torch
    from delayed_speech_synthesis_b200 import transforms as TR
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, c_raw = 1000, 30_007, 129
    raw = O.synthetic_ecog(T, c_raw, fs, "ecog", session=53)
    pre_o, pre = _chain(c_raw, 4)
    mapped = raw
    for f in pre_o:
        mapped = f(mapped)
    want = _oracle_offline(mapped, fs)
    eng = HgaEngine(fs, 64, precision="f64", device=0)
    got = eng.extract_offline_mapped(torch.from_numpy(raw).cuda(), TR.ChannelMap(TR.fold(pre, c_raw), 0)).cpu().numpy()
    assert not eng.last_fused and np.abs(got - want).max() <= TOL_F64


def test_extract_trials_takes_the_fused_path():
    from delayed_speech_synthesis_b200 import corpus
    fs, c_raw = 1000, 129
    raw = O.synthetic_ecog(20_000, c_raw, fs, "ecog", session=54)
    pre_o, pre = _chain(c_raw, 5)
    trials = [("A", 100, 6_300), ("B", 6_280, 12_000), ("C", 12_500, 19_900)]
    feats = []
    for _, a, b in trials:
        ex = O.HighGammaExtractor(fs, 64, pre_transforms=pre_o)
        feats.append(ex.extract_features(raw[a:int(b + 0.04 * fs)]))
    want = np.concatenate(feats)
    got = corpus.extract_trials(raw, trials, fs, 64, pre, precision="f32")
    assert got.shape == want.shape and np.abs(got - want).max() <= TOL_F32


def test_fused_many_short_segments_several_items_per_warp():
    """More (segment, channel group) items than the device runs at a time: every mapper / stage pair walks through several
    items, and the hand-off barriers carry their phase from one item to the next."""
    import torch
    from delayed_speech_synthesis_b200 import transforms as TR
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, c_raw, n_seg, seg = 1000, 129, 700, 333
    raw = O.synthetic_ecog(n_seg * seg, c_raw, fs, "ecog", session=55)
    pre_o, pre = _chain(c_raw, 6)
    mapped = raw
    for f in pre_o:
        mapped = f(mapped)
    segs = [(i * seg, i * seg + seg - (i % 7)) for i in range(n_seg)]
    want = _oracle_offline(mapped, fs, segments=segs)
    eng = HgaEngine(fs, 64, precision="f32", device=0)
    got = eng.extract_offline_mapped(torch.from_numpy(raw).cuda(), TR.ChannelMap(TR.fold(pre, c_raw), 0), segs).cpu().numpy()
    assert eng.last_fused and got.shape == want.shape
    assert np.abs(got - want).max() <= TOL_F32


def test_fused_padded_rows_last_rows_and_two_channel_groups_with_means():
    """Raw rows with a leading dimension wider than the map reads (row-by-row copies instead of one span), a segment that
    ends on the recording's last row (the mappers' look-ahead is clamped there), and a map that keeps both grids after the
    common average: 128 output channels = two items per chunk, each reading means of both groups' tables."""
    import torch
    from delayed_speech_synthesis_b200 import transforms as TR
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T, c_raw = 1000, 33_333, 130
    raw = O.synthetic_ecog(T, c_raw, fs, "ecog", session=56) + np.linspace(300.0, -200.0, c_raw)[None, :]
    layout = np.arange(128) + 1
    grids = [layout[:64].reshape(8, 8), np.flip(layout[64:].reshape(8, 8), axis=1)]
    bad = [1, 64, 65, 128]
    both = np.arange(2, 130)
    appl, comp = O.car_masks(bad, grids, layout)
    mapped = O.common_average_reference(O.select_columns(raw, both), appl, comp)
    pre = [TR.ChannelSelection(both), TR.CommonAverageReferencing(bad, grids, layout)]
    segs = [(0, 20_000), (20_000, T)]
    want = _oracle_offline(mapped, fs, segments=segs)
    wide = torch.full((T, c_raw + 6), float("nan"), dtype=torch.float64, device="cuda")
    wide[:, :c_raw] = torch.from_numpy(raw).cuda()
    eng = HgaEngine(fs, 128, precision="f32", device=0)
    got = eng.extract_offline_mapped(wide[:, :c_raw], TR.ChannelMap(TR.fold(pre, c_raw), 0), segs).cpu().numpy()
    assert eng.last_fused and np.isfinite(got).all()
    assert np.abs(got - want).max() <= TOL_F32
