"""`hga_stats_accumulate_static` (include/hga_b200.h): the baseline-statistics accumulation with its block partials in a
static device array — the entry point for callers that cannot call cudaMalloc.  Called through the C ABI (ctypes), checked
against the existing `hga_stats_accumulate` on two calls in a row (the second one smaller, so the shared scratch still holds
partials of the first), against numpy (rtol 1e-12, the tolerance of the statistics tests), and for the size limit.
Collected last on purpose: nothing runs after it."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-12
HGA_ERR_INVALID = 1           # include/hga_b200.h


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def test_static_scratch_accumulate_matches_the_allocating_entry_point():
    import torch
    from delayed_speech_synthesis_b200 import _lib
    from delayed_speech_synthesis_b200._lib import check

    lib = _lib.load()
    dev = 0
    torch.cuda.set_device(dev)
    g = torch.Generator(device="cuda").manual_seed(20241104)
    Cn = 128
    f1 = torch.randn((35_999, Cn), dtype=torch.float64, device="cuda", generator=g) * 0.7 + 3.0
    f2 = torch.randn((1_237, Cn + 5), dtype=torch.float64, device="cuda", generator=g)[:, :Cn] * 2.0 - 1.0   # padded rows: ld > C
    acc_ref = torch.zeros((3, Cn), dtype=torch.float64, device="cuda")
    acc_new = torch.zeros((3, Cn), dtype=torch.float64, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for f in (f1, f2):                                   # two calls in a row on the shared scratch
        check(lib.hga_stats_accumulate(dev, _ptr(f), f.shape[0], Cn, f.stride(0), _ptr(acc_ref), stream))
        check(lib.hga_stats_accumulate_static(dev, _ptr(f), f.shape[0], Cn, f.stride(0), _ptr(acc_new), stream))
        torch.cuda.synchronize()
        a, b = acc_ref.cpu().numpy(), acc_new.cpu().numpy()
        assert np.allclose(b, a, rtol=RTOL, atol=0.0)
        assert np.array_equal(a, b)                      # same tiling, same merge order: the same bits
    mu = torch.empty(Cn, dtype=torch.float64, device="cuda")
    sg = torch.empty_like(mu)
    check(lib.hga_stats_finalize(dev, _ptr(acc_new), Cn, _ptr(mu), _ptr(sg), stream))
    torch.cuda.synchronize()
    allf = np.concatenate([f1.cpu().numpy(), f2.cpu().numpy()])
    assert np.allclose(mu.cpu().numpy(), allf.mean(0), rtol=RTOL)
    assert np.allclose(sg.cpu().numpy(), allf.std(0), rtol=1e-11)
    assert acc_new[0].cpu().numpy().tolist() == [float(len(allf))] * Cn

    # a second pair after unrelated work, other widths (ragged channel groups, one channel)
    for Cw, F in ((100, 4_099), (1, 777), (1024, 300)):
        f = torch.randn((F, Cw), dtype=torch.float64, device="cuda", generator=g)
        a = torch.zeros((3, Cw), dtype=torch.float64, device="cuda")
        b = torch.zeros((3, Cw), dtype=torch.float64, device="cuda")
        check(lib.hga_stats_accumulate(dev, _ptr(f), F, Cw, f.stride(0), _ptr(a), stream))
        check(lib.hga_stats_accumulate_static(dev, _ptr(f), F, Cw, f.stride(0), _ptr(b), stream))
        torch.cuda.synchronize()
        assert np.array_equal(a.cpu().numpy(), b.cpu().numpy())
        assert np.allclose(b[1].cpu().numpy(), f.cpu().numpy().mean(0), rtol=1e-9, atol=1e-12)


def test_static_scratch_rejects_shapes_beyond_its_size():
    import torch
    from delayed_speech_synthesis_b200 import _lib

    lib = _lib.load()
    torch.cuda.set_device(0)
    dummy = torch.zeros(8, dtype=torch.float64, device="cuda")     # never read: the size check comes first
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.hga_stats_accumulate_static(0, _ptr(dummy), 10, 50_000, 50_000, _ptr(dummy), stream)
    assert rc == HGA_ERR_INVALID
    msg = lib.hga_last_error().decode()
    assert "static scratch" in msg and "50000" in msg
    # F == 0 is a no-op, as in the allocating entry point
    assert lib.hga_stats_accumulate_static(0, _ptr(dummy), 0, 2, 2, _ptr(dummy), stream) == 0
