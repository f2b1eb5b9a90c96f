"""Memory-safety checks of our own (compute-sanitizer is closed on this pool): every kernel family runs with
 * guard bands around each device buffer the library allocates (`hga_debug_guards`: 64 KB of a fill pattern in front of and
   behind every scratch / state buffer, read back by `hga_debug_check_guards`),
 * the caller's input embedded in a larger NaN-filled buffer (a read outside the rows / columns handed over would put a
   NaN into a feature) and the caller's output embedded in a canary-filled buffer (a write outside it would change a
   canary; a frame left unwritten would keep the canary),
and the results must still match the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import hga_oracle as O  # noqa: E402

from test_gpu_parity import TOL_F32, TOL_F64, TOL_MIX, _oracle_offline  # noqa: E402

CANARY = -777.25


@pytest.fixture(scope="module")
def guarded():
    from delayed_speech_synthesis_b200 import _lib
    lib = _lib.load()
    lib.hga_debug_guards(1)
    yield lib
    lib.hga_debug_guards(0)


def _embedded(torch, x, pad_rows, pad_cols):
    """x inside a NaN frame: returns (view with the values of x, backing buffer)."""
    T, C = x.shape
    big = torch.full((T + 2 * pad_rows, C + pad_cols), float("nan"), dtype=torch.float64, device="cuda")
    view = big[pad_rows:pad_rows + T, :C]
    view.copy_(torch.from_numpy(x))
    return view, big


@pytest.mark.parametrize("precision,tol", [("f64", TOL_F64), ("f32", TOL_F32), ("mixed", TOL_MIX), ("exact", 4e-15)])
@pytest.mark.parametrize("C,pad_cols", [(128, 0), (100, 4), (33, 1)])
def test_offline_calls_stay_inside_their_buffers(guarded, precision, tol, C, pad_cols):
    import torch
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, T = 1000, 41_237                                  # several carried chunks, a ragged last hop block
    x = O.synthetic_ecog(T, C, fs, "ecog", session=31)
    rng = np.random.default_rng(5)
    mu, sigma = rng.normal(3, 0.1, C), rng.uniform(0.3, 0.6, C)
    segs = [(0, 9_000), (9_000, 9_050), (9_100, 30_007), (30_500, T)]
    want = _oracle_offline(x, fs, mu, sigma, segs)
    xv, backing = _embedded(torch, x, 64, pad_cols)
    eng = HgaEngine(fs, C, precision=precision, device=0)
    F = want.shape[0]
    out_big = torch.full((F + 16, C + 3), CANARY, dtype=torch.float64, device="cuda")
    out = out_big[8:8 + F, :C]
    eng.extract_offline(xv, segs, torch.from_numpy(mu).cuda(), torch.from_numpy(sigma).cuda(), out=out)
    assert guarded.hga_debug_guarded_buffers() > 0
    assert guarded.hga_debug_check_guards() == 0
    got = out.cpu().numpy()
    assert np.isfinite(got).all() and not (got == CANARY).any()
    scale = 1.0 / sigma.min()
    assert np.abs(got - want).max() <= tol * max(1.0, scale)
    ob = out_big.cpu().numpy()
    ob[8:8 + F, :C] = CANARY
    assert (ob == CANARY).all()                            # nothing outside the output view was written
    assert torch.isnan(backing[:64]).all() and torch.isnan(backing[64 + T:]).all()   # the input was not written either
    del eng


def test_host_path_statistics_and_streaming_stay_inside_their_buffers(guarded):
    import torch
    from delayed_speech_synthesis_b200 import ops
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C, T = 1000, 64, 30_011
    x = O.synthetic_ecog(T, C, fs, "white", session=32)
    want = _oracle_offline(x, fs)
    for precision, tol in (("f64", TOL_F64), ("f32", TOL_F32)):
        eng = HgaEngine(fs, C, precision=precision, device=0)
        got = eng.extract_offline(x)                       # pageable host buffers: staging, slabs, partial launches
        assert np.abs(got - want).max() <= tol
        assert guarded.hga_debug_check_guards() == 0
        del eng
    feats = torch.from_numpy(want).cuda()
    st = ops.ChannelStats(C, 0).update(feats[:1000]).update(feats[1000:])
    mu, sg = st.finalize()
    assert np.allclose(mu, want.mean(0), rtol=1e-12) and np.allclose(sg, want.std(0), rtol=1e-10)
    eng = HgaEngine(fs, C, precision="exact", device=0)
    frames = [eng.push(x[i * 10:(i + 1) * 10]) for i in range(200)]
    n = sum(len(f) for f in frames)
    assert n == 200                                        # 10-row pushes: the first is zero-padded to one window, then one frame per push
    assert guarded.hga_debug_check_guards() == 0
    del eng
