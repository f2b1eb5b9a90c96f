"""Channel-axis transforms either side of the HGA path.

CPU part: the oracle's numpy restatements against the golden vectors produced by the reference's own classes
(tests/golden/make_golden.py, /root/reference/local/common.py:16-58,220-343), and the host-side folding of a chain into
one (sel, grp, alpha, beta, groups) map, evaluated in numpy.
GPU part: the channel-map kernel through the C ABI against the same vectors and against the oracle on seeded inputs.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import pytest

from oracle import hga_oracle as O
from delayed_speech_synthesis_b200 import transforms as TR

G = np.load(Path(__file__).parent / "golden" / "transforms.npz")
TOL = 1e-12   # max-abs on ~100 uV data: only the summation order of the means differs (numpy pairwise vs lane-strided)


def _grids():
    return [G["speech_grid"], G["motor_grid"]]


def eval_spec(spec: TR.ChannelMapSpec, x: np.ndarray) -> np.ndarray:
    """The map hga_chanmap_apply computes, in numpy (host logic check, no GPU)."""
    out = np.zeros((len(x), spec.c_out))
    for j in range(spec.c_out):
        if spec.sel[j] >= 0:
            out[:, j] = spec.alpha[j] * x[:, spec.sel[j]]
        if spec.grp[j] >= 0:
            out[:, j] += spec.beta[j] * x[:, spec.groups[spec.grp[j]]].mean(axis=1)
    return out


# ------------------------------------------------------------------------------------------------------ oracle pins
def test_oracle_selection_and_car_match_reference_classes():
    x = G["x"]
    a = O.select_columns(x, G["both_mapping"])
    assert np.array_equal(a, G["after_both"])
    appl, comp = O.car_masks(G["exclude"], _grids(), G["layout"])
    assert np.array_equal(np.array(appl), G["car_appl"]) and np.array_equal(np.array(comp), G["car_comp"])
    b = O.common_average_reference(a, appl, comp)
    assert np.array_equal(b, G["after_car"])
    assert np.array_equal(O.select_columns(b, G["speech_mapping"]), G["after_speech"])


def test_oracle_bad_channel_correction_matches_reference_class():
    patches = O.bad_channel_patches(G["bad"], _grids(), G["layout"])
    assert [int(p[0][0]) for p in patches] == list(G["bcc_loc"])
    for p, want in zip(patches, G["bcc_neigh"]):
        assert list(p[1]) == list(want[want >= 0])
    assert np.array_equal(O.bad_channel_correction(G["f"], patches), G["after_bcc"])


# ------------------------------------------------------------------------------------------------------ host folding
def _chain():
    both = TR.ChannelSelection(G["both_mapping"])
    car = TR.CommonAverageReferencing(list(G["exclude"]), _grids(), G["layout"])
    speech = TR.ChannelSelection(G["speech_mapping"])
    return both, car, speech


def test_mirrored_classes_build_the_reference_tables():
    _, car, _ = _chain()
    assert np.array_equal(np.array(car.selection_masks_application), G["car_appl"])
    assert np.array_equal(np.array(car.selection_masks_computation), G["car_comp"])
    bcc = TR.BadChannelCorrection(list(G["bad"]), _grids(), G["layout"])
    assert [int(p[0][0]) for p in bcc.patches] == list(G["bcc_loc"])
    for p, want in zip(bcc.patches, G["bcc_neigh"]):
        assert list(p[1]) == list(want[want >= 0])
    assert len(bcc) == 7 and "Correcting 7 bad channels" in repr(bcc)


def test_fold_chain_is_one_map_in_raw_columns():
    both, car, speech = _chain()
    x = G["x"]
    for chain, want in (([both], G["after_both"]), ([both, car], G["after_car"]), ([both, car, speech], G["after_speech"])):
        spec = TR.fold(chain, 128)
        assert spec is not None and spec.c_out == want.shape[1]
        assert np.abs(eval_spec(spec, x) - want).max() <= TOL
    spec = TR.fold([TR.BadChannelCorrection(list(G["bad"]), _grids(), G["layout"])], 128)
    assert np.abs(eval_spec(spec, G["f"]) - G["after_bcc"]).max() <= TOL
    assert (spec.sel[G["bcc_loc"]] == -1).all()      # replaced columns have no direct term


def test_fold_refuses_what_one_map_cannot_express():
    both, car, speech = _chain()
    assert TR.fold([both, car, car], 128) is None                      # two averaging stages
    assert TR.fold([lambda d: d * 2.0], 128) is None                   # unknown callable
    assert TR.fold([both], 64) is None                                 # index out of range for the input width
    assert TR.fold(None, 128) is None and TR.fold([], 128) is None
    many_bad = TR.BadChannelCorrection(list(range(1, 128, 2)) + [2, 4], _grids(), G["layout"])   # 66 patches > 64 groups
    assert len(many_bad) == 66 and TR.fold([many_bad], 128) is None


def test_fold_recognises_reference_objects_by_attribute():
    class SelectElectrodesFromBothGrids:                               # shape of local/common.py:16-32
        def __init__(self):
            self.grid_mapping = list(G["both_mapping"])

    class CommonAverageReferencing:                                    # shape of local/common.py:308-343
        selection_masks_application = list(G["car_appl"])
        selection_masks_computation = list(G["car_comp"])

    spec = TR.fold([SelectElectrodesFromBothGrids(), CommonAverageReferencing()], 128)
    assert spec is not None and np.abs(eval_spec(spec, G["x"]) - G["after_car"]).max() <= TOL


# ------------------------------------------------------------------------------------------------------ GPU, C ABI
@pytest.mark.gpu
def test_chanmap_kernel_matches_reference_vectors():
    both, car, speech = _chain()
    x = G["x"]
    assert np.array_equal(both(x), G["after_both"])                     # pure gather: bit-exact
    assert np.abs(car(G["after_both"]) - G["after_car"]).max() <= TOL
    m = TR.ChannelMap(TR.fold([both, car, speech], 128))
    assert np.abs(m.apply(x) - G["after_speech"]).max() <= TOL
    bcc = TR.BadChannelCorrection(list(G["bad"]), _grids(), G["layout"])
    got = bcc(G["f"])
    assert np.abs(got - G["after_bcc"]).max() <= TOL
    keep = np.setdiff1d(np.arange(128), G["bcc_loc"])
    assert np.array_equal(got[:, keep], G["f"][:, keep])


@pytest.mark.gpu
@pytest.mark.parametrize("T,c_in", [(1, 128), (7, 40), (4099, 128), (1000, 1024)])
def test_chanmap_kernel_vs_oracle_seeded(T, c_in):
    import torch

    rng = np.random.Generator(np.random.Philox(key=[20241104, T + c_in]))
    x = rng.standard_normal((T, c_in)) * 30.0
    perm = rng.permutation(c_in)
    n_grid = c_in // 2
    layout = np.arange(c_in) + 1
    grids = [layout[:n_grid].reshape(-1, 4), layout[n_grid:2 * n_grid].reshape(-1, 4)]
    excl = [int(layout[1]), int(layout[n_grid + 2])]
    chain = [TR.ChannelSelection(perm), TR.CommonAverageReferencing(excl, grids, layout), TR.ChannelSelection(np.arange(0, c_in, 2))]
    appl, comp = O.car_masks(excl, grids, layout)
    want = O.select_columns(O.common_average_reference(O.select_columns(x, perm), appl, comp), np.arange(0, c_in, 2))
    m = TR.ChannelMap(TR.fold(chain, c_in))
    assert np.abs(m.apply(x) - want).max() <= TOL                       # host buffers
    xd = torch.from_numpy(x).cuda()
    assert np.abs(m.apply(xd).cpu().numpy() - want).max() <= TOL        # device buffers
    pad = torch.zeros((T, c_in + 8), dtype=torch.float64, device="cuda")
    pad[:, :c_in] = xd
    assert torch.equal(m.apply(pad[:, :c_in]), m.apply(xd))             # ldx > c_in


@pytest.mark.gpu
def test_chanmap_errors():
    with pytest.raises(Exception):
        TR.ChannelMap(TR.ChannelMapSpec(8, np.array([9], np.int32), np.array([-1], np.int32), np.ones(1), np.zeros(1), []))
    m = TR.ChannelMap(TR.ChannelMapSpec.identity(8))
    with pytest.raises(IndexError):
        m.apply(np.zeros((4, 9)))
    assert m.apply(np.zeros((0, 8))).shape == (0, 8)


@pytest.mark.gpu
def test_extractor_with_reference_transform_chain():
    """prepare_corpus.py:152-175 wiring: select -> CAR -> select before the filters, bad-channel patch after framing."""
    from delayed_speech_synthesis_b200.units import HighGammaExtractor

    fs, T = 1000, 3000
    x = O.synthetic_ecog(T, 128, fs, "ecog")
    both, car, speech = _chain()
    appl, comp = O.car_masks(G["exclude"], _grids(), G["layout"])
    bad = [19, 38]
    layout64 = np.arange(64) + 1          # the 64 selected columns, numbered as the speech grid numbers them
    patches = O.bad_channel_patches(bad, [G["speech_grid"]], layout64)
    pre_o = [lambda d: O.select_columns(d, G["both_mapping"]), lambda d: O.common_average_reference(d, appl, comp),
             lambda d: O.select_columns(d, G["speech_mapping"])]
    ref = O.HighGammaExtractor(fs, 64, pre_transforms=pre_o, post_transforms=[lambda f: O.bad_channel_correction(f, patches)])
    want = ref.extract_features(x)
    bcc = TR.BadChannelCorrection(bad, [G["speech_grid"]], layout64)
    ex = HighGammaExtractor(fs, 64, pre_transforms=[both, car, speech], post_transforms=[bcc])
    got = ex.extract_features(x)
    assert got.shape == want.shape and isinstance(got, np.ndarray)
    assert np.abs(got - want).max() <= 1e-9
    # streaming packets through the same chain: the folded map runs per packet
    ex2 = HighGammaExtractor(fs, 64, pre_transforms=[both, car, speech], post_transforms=[bcc])
    ref2 = O.HighGammaExtractor(fs, 64, pre_transforms=pre_o, post_transforms=[lambda f: O.bad_channel_correction(f, patches)])
    for t in range(0, 200, 10):
        a, b = ex2.extract_features(x[t:t + 10]), ref2.extract_features(x[t:t + 10])
        assert a.shape == b.shape and np.abs(a - b).max() <= 1e-11


@pytest.mark.gpu
def test_bad_channel_correction_many_patches():
    """Every second channel bad (64 patches = the kernel's group table full): all group means reduced in four passes."""
    bad = list(range(1, 128, 2))
    bcc = TR.BadChannelCorrection(bad, _grids(), G["layout"])
    want = O.bad_channel_correction(G["f"], O.bad_channel_patches(bad, _grids(), G["layout"]))
    assert np.abs(bcc(G["f"]) - want).max() <= TOL
