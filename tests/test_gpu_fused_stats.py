"""Baseline statistics out of the extraction kernels (`hga_extract_offline_stats`): the day-baseline flow of
/root/reference/prepare_corpus.py:190-199 (features of the baseline recording's trials, then np.mean / np.std over all
frames) in one pass.  mu within rtol 1e-12 and sigma within rtol 1e-10 of numpy over the features the same call returned
(the tolerance of the statistics tests), the features bit-identical to the plain extraction."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import hga_oracle as O  # noqa: E402


@pytest.mark.parametrize("precision,C,fused", [("f32", 128, True), ("f32", 100, True), ("f32", 34, True), ("f64", 128, False), ("f32", 33, False), ("mixed", 64, False),
                                               ("exact", 16, False)])
def test_statistics_leave_the_extraction_kernels(precision, C, fused):
    import torch
    from delayed_speech_synthesis_b200 import ops
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T = 1000, 150_011
    x = torch.from_numpy(O.synthetic_ecog(T, C, fs, "ecog", session=61)).cuda()
    segs = [(0, 90_000), (90_000, 90_050), (90_100, 90_777), (91_000, T)]       # carried chunks, a one-frame trial, a ragged tail
    eng = HgaEngine(fs, C, precision=precision, device=0)
    st = ops.ChannelStats(C, 0)
    feats = eng.extract_offline_stats(x, st, segs)
    assert eng.last_fused == fused
    plain = eng.extract_offline(x, segs)
    assert torch.equal(feats, plain)                                             # the statistics kernels write the same frames
    f = feats.cpu().numpy()
    acc = st.acc.cpu().numpy()
    assert np.array_equal(acc[0], np.full(C, float(len(f))))
    mu, sg = st.finalize()
    assert np.allclose(mu, f.mean(0), rtol=1e-12, atol=1e-13)
    assert np.allclose(sg, f.std(0), rtol=1e-10)
    # a second recording accumulates on top (the per-day loop over baseline trials), against the two-pass accumulator
    x2 = torch.from_numpy(O.synthetic_ecog(40_000, C, fs, "white", session=62)).cuda()
    f2 = eng.extract_offline_stats(x2, st)
    ref = ops.ChannelStats(C, 0).update(feats).update(f2)
    m1, s1 = st.finalize()
    m2, s2 = ref.finalize()
    assert np.allclose(m1, m2, rtol=1e-12, atol=1e-13) and np.allclose(s1, s2, rtol=1e-10)
    allf = np.concatenate([f, f2.cpu().numpy()])
    assert np.allclose(m1, allf.mean(0), rtol=1e-12, atol=1e-13) and np.allclose(s1, allf.std(0), rtol=1e-10)


def test_statistics_of_z_scored_output_and_single_chunk():
    """With mu / sigma the stored (z-scored) values are what is accumulated; a recording shorter than the carry is one chunk
    (no fix-up kernels run)."""
    import torch
    from delayed_speech_synthesis_b200 import ops
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C, T = 1000, 64, 310
    x = torch.from_numpy(O.synthetic_ecog(T, C, fs, "white", session=63)).cuda()
    rng = np.random.default_rng(1)
    mu, sigma = rng.normal(3, 0.2, C), rng.uniform(0.3, 0.6, C)
    eng = HgaEngine(fs, C, precision="f32", device=0)
    st = ops.ChannelStats(C, 0)
    z = eng.extract_offline_stats(x, st, None, mu, sigma).cpu().numpy()
    assert eng.last_fused and eng.last_plan()["carried"] == 0
    m, s = st.finalize()
    assert np.allclose(m, z.mean(0), rtol=1e-11, atol=1e-12) and np.allclose(s, z.std(0), rtol=1e-10)


def test_statistics_at_full_k2_size_with_channel_gains_over_five_decades():
    """K2 (128 ch x 3600 s @ 1 kHz), per-channel gains 1e-2 .. 1e3 so that the channel means of the log-power features spread
    from -4.6 (the 0.01 floor inside the logarithm) to +12 around the kernels' fixed shift of 3: mu / sigma against a float64 two-pass mean / std of the features the
    same call returned (computed on the device), and the frame count."""
    import torch
    from delayed_speech_synthesis_b200 import ops
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C, T = 1000, 128, 3_600_000
    g = torch.Generator(device="cuda").manual_seed(64)
    x = torch.randn((T, C), dtype=torch.float64, device="cuda", generator=g)
    x *= torch.logspace(-2, 3, C, dtype=torch.float64, device="cuda")
    x += 5000.0                                                                    # amplifier offset: out of band, removed by the band-pass
    eng = HgaEngine(fs, C, precision="f32", device=0)
    st = ops.ChannelStats(C, 0)
    feats = eng.extract_offline_stats(x, st)
    assert eng.last_fused and eng.last_plan()["carried"] > 100
    del x
    mu, sg = st.finalize(as_numpy=False)
    m_ref = feats.mean(0)
    s_ref = (feats - m_ref).square().mean(0).sqrt()                                # np.std: population form
    assert float(m_ref.max() - m_ref.min()) > 15.0
    print('sigma range', float(s_ref.min()), float(s_ref.max()), 'max rel err', float(((sg - s_ref).abs() / s_ref).max()))
    assert st.acc[0].cpu().numpy().tolist() == [float(feats.shape[0])] * C
    assert torch.allclose(mu, m_ref, rtol=1e-12, atol=1e-12)
    assert torch.allclose(sg, s_ref, rtol=1e-9)
