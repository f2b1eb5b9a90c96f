"""CPU: the C-ABI library loads and exports every symbol include/hga_b200.h declares (no compute calls without a GPU),
it refuses to run without a device (no CPU fallback), and the host-side logic (filter design, window geometry, frame
buffer, ezmsg stand-in, sharding) behaves like the reference's."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

from oracle import hga_oracle as O

ROOT = Path(__file__).resolve().parents[1]


def _declared_symbols():
    text = (ROOT / "include" / "hga_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hga_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from delayed_speech_synthesis_b200 import _lib
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(names)
    assert lib.hga_abi_version() == 1


def test_config_struct_layout_matches_header():
    from delayed_speech_synthesis_b200._lib import Config
    # 6 int32 + double + 2 int32 + (48+48+16+16) doubles + 2 int32 + double + 2 int32
    assert ctypes.sizeof(Config) == 24 + 8 + 8 + 128 * 8 + 8 + 8 + 8
    assert Config.eps.offset == 24 and Config.sos_bp.offset == 40 and Config.precision.offset == 40 + 128 * 8


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from delayed_speech_synthesis_b200 import HgaError, _lib
    from delayed_speech_synthesis_b200.units import HighGammaExtractor
    from delayed_speech_synthesis_b200.hga_optimized import compute_log_power_features
    assert _lib.load().hga_device_count() == 0
    with pytest.raises(HgaError):
        HighGammaExtractor(1000, 8)
    with pytest.raises(HgaError):
        compute_log_power_features(np.zeros((100, 2)), 1000, 0.05, 0.01)
    # and the C entry points themselves refuse (status 3), they do not compute on the host
    lib = _lib.load()
    sos = np.ascontiguousarray(O.design_hga_filters(1000)[0])
    assert lib.hga_sosfilt(0, sos.ctypes.data, 8, 1, 10, 1, 1, 1, 1, 1, None) == 3
    assert b"no CPU fallback" in lib.hga_last_error()


def test_product_never_imports_oracle():
    for p in (ROOT / "delayed_speech_synthesis_b200").rglob("*.py"):
        assert not re.search(r"^\s*(from|import)\s+oracle\b", p.read_text(), flags=re.M), p


def test_filter_design_matches_oracle_and_appendix_b():
    from delayed_speech_synthesis_b200 import filters
    for fs in (1000, 2000):
        bp, nt = filters.design_hga_filters(fs)
        obp, ont = O.design_hga_filters(fs)
        assert np.array_equal(bp, obp) and np.array_equal(nt, ont)
        assert np.array_equal(filters.tiled_zi(bp, 5), O.tiled_zi(bp, 5))
    bp, nt = filters.design_hga_filters(1000)
    assert bp.shape == nt.shape == (8, 6)
    assert abs(bp[0, 0] - 2.3959644103776190e-05) < 1e-18 and abs(bp[7, 4] + 1.7412122991087324) < 1e-13
    assert abs(nt[0, 1] + 1.367092559893505) < 1e-13 and abs(nt[7, 5] - 0.9951733587893605) < 1e-13
    # the structure the fused kernels rely on: b0 == b2 in every section, gain only in section 0
    for sos in (bp, nt):
        assert np.array_equal(sos[:, 0], sos[:, 2]) and np.all(sos[1:, 0] == 1.0) and np.all(sos[:, 3] == 1.0)


def test_window_geometry_and_indices_match_oracle():
    from delayed_speech_synthesis_b200 import filters
    assert filters.window_geometry(1000, 0.05, 0.01) == (50, 10)
    assert filters.window_geometry(2000, 0.05, 0.01) == (100, 20)
    for rows, sr, wl, ws in ((1000, 1000, 0.05, 0.01), (700, 1000, 0.05, 0.0125), (49, 1000, 0.05, 0.01), (333, 1000, 0.032, 0.008)):
        a, b = filters.window_indices(rows, sr, wl, ws)
        oa, ob = O.window_indices(rows, sr, wl, ws)
        assert np.array_equal(a, oa) and np.array_equal(b, ob)
    s, e = filters.window_indices(700, 1000, 0.05, 0.0125)
    assert not filters.windows_are_regular(s, e, 50, 12, 700)
    s, e = filters.window_indices(60000, 1000, 0.05, 0.01)
    assert filters.windows_are_regular(s, e, 50, 10, 60000)


def test_host_frame_buffer_matches_compiled_reference_class(golden_dir):
    from delayed_speech_synthesis_b200.hga_optimized import WarmStartFrameBuffer
    for name in ("framebuffer_a", "framebuffer_b", "framebuffer_c"):
        g = np.load(golden_dir / f"{name}.npz")
        fb = WarmStartFrameBuffer(0.05, 0.01, 1000, g["in0"].shape[1])
        for i in range(len(g["chunks"])):
            assert np.array_equal(fb.insert(g[f"in{i}"]), g[f"out{i}"]), (name, i)
        fb.reset()
        assert fb.first_frame and not fb.remainder_data.any()
    with pytest.raises(ValueError):
        WarmStartFrameBuffer(0.05, 0.01, 1000, 2).insert(np.zeros((10, 2), dtype=np.float32))


def test_ezmsg_standin_and_unit_declaration():
    from delayed_speech_synthesis_b200.ezmsg_compat import TimeSeriesMessage, ez
    from delayed_speech_synthesis_b200.units import HighGammaActivity, HighGammaActivitySettings, HighGammaActivityState
    s = HighGammaActivitySettings(fs=1000, nb_electrodes=64)
    assert (s.window_length, s.window_shift, s.l_freq, s.h_freq, s.pre_transforms, s.post_transforms) == (0.05, 0.01, 70, 170, None, None)
    with pytest.raises(TypeError):
        HighGammaActivitySettings()                                  # fs, nb_electrodes are required
    u = HighGammaActivity(s)
    assert isinstance(u.STATE, HighGammaActivityState) and u.STATE.hg_extractor is None
    assert isinstance(HighGammaActivity.INPUT, ez.InputStream) and isinstance(HighGammaActivity.OUTPUT, ez.OutputStream)
    m = TimeSeriesMessage(data=np.zeros((4, 2)), fs=1000.0)
    from dataclasses import replace
    assert replace(m, fs=100.0).fs == 100.0 and m.time_dim == 0


def test_session_sharding():
    from delayed_speech_synthesis_b200.corpus import day_slots, shard_sessions, shard_weighted
    w = [2 if s % 8 == 0 else 1 for s in range(64)]        # k5: the first session of a day is its baseline (extracted twice)
    for world in (1, 2, 4, 8):
        parts = [shard_weighted(w, r, world) for r in range(world)]
        assert sorted(s for p in parts for s in p) == list(range(64))
        assert {sum(w[s] for s in p) for p in parts} == {72 // world}
    for world in (1, 2, 4, 8):
        got = sorted(s for r in range(world) for s in shard_sessions(64, r, world))
        assert got == list(range(64))
        assert max(len(shard_sessions(64, r, world)) for r in range(world)) == 64 // world
    assert shard_sessions(5, 1, 2) == [1, 3]
    with pytest.raises(ValueError):
        shard_sessions(4, 2, 2)
    assert day_slots([3, 1, 3, 2]) == {1: 0, 2: 1, 3: 2}
