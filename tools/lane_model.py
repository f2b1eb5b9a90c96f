"""NumPy model of the lane kernels' algorithm (hga_kernels_lane.cuh): the kernel's own section order and forms, the fresh
state in that basis, zero-state time-chunks, the chunk-to-chunk state carry (A^Lc propagation) and the staged zero-input
fix-up of each chunk's head.  Design aid + numerical check of the scheme against the CPU oracle (tools only; the product
never imports it).  Run:  python tools/lane_model.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from oracle import hga_oracle as O  # noqa: E402


def lane_operands(fs):
    """Kernel-order operands: sections 0..7 'A' (numerator [1, 0, -1], negated state), 8..15 'U' (numerator [1, b1, 1])."""
    hg, fh = O.design_hga_filters(fs)
    plus = [s for s in range(8) if abs(hg[s, 1] / hg[s, 0] - 2) < 1e-12]
    minus = [s for s in range(8) if abs(hg[s, 1] / hg[s, 0] + 2) < 1e-12]
    assert len(plus) == 4 and len(minus) == 4
    G = hg[0, 0] * fh[0, 0]
    c1 = np.zeros(16); c2 = np.zeros(16); cb = np.zeros(16)
    for s in range(8):        # denominators keep the caller's order; numerators re-paired (1+z^-1)(1-z^-1)
        c1[s], c2[s] = hg[s, 4], hg[s, 5]
    for s in range(8):
        c1[8 + s], c2[8 + s], cb[8 + s] = -fh[s, 4], -fh[s, 5], fh[s, 1] / fh[s, 0]
    # fresh state: band-pass = steady state of the history x == 1 (only section 0 is excited: DC gain 0 after it);
    # notch = sosfilt_zi of the notch divided by the gains in front (the kernel runs on the gain-free signal)
    f0 = np.zeros(16); f1 = np.zeros(16)
    f0[0] = 1.0; f1[0] = 1.0   # v = w = 1 : y = x - v = 0, v' = a1*0 + w = 1, w' = a2*0 + x = 1
    zi = O.sosfilt_zi(fh)
    g = hg[0, 0]
    for s in range(8):
        g = g * fh[s, 0]
        f0[8 + s] = zi[s, 0] / g
        f1[8 + s] = zi[s, 1] / g
    return dict(c1=c1, c2=c2, cb=cb, f0=f0, f1=f1, G=G, hg=hg, fh=fh)


def run_cascade(op, x, z0, z1, first=0, dtype=np.float64):
    """x: (T, C) or None for zero input of T rows; sections [first, 16).  Returns y (T, C); z0/z1 (16, C) updated in place."""
    c1, c2, cb = (op[k].astype(dtype) for k in ("c1", "c2", "cb"))
    T = x.shape[0]
    y = np.empty((T, x.shape[1]), dtype)
    for t in range(T):
        u = x[t].astype(dtype)
        for s in range(first, 16):
            if s < 8:
                v = u - z0[s]
                z0[s] = c1[s] * v + z1[s]
                z1[s] = c2[s] * v + u
            else:
                v = u + z0[s]
                z0[s] = c1[s] * v + (cb[s] * u + z1[s])
                z1[s] = c2[s] * v + u
            u = v
        y[t] = u
    return y


def features_from_y(y, op, L, hop, eps=0.01):
    T = y.shape[0]
    F = (T - L) // hop + 1
    nb = T // hop
    bs = (y[: nb * hop].astype(np.float64) ** 2).reshape(nb, hop, -1).sum(1)
    out = np.empty((F, y.shape[1]))
    for w in range(F):
        out[w] = np.log(bs[w:w + L // hop].sum(0) * (op["G"] ** 2 / L) + eps)
    return out


def transition_matrix(op, n):
    """A^n in the kernel's state basis (z0[0..15], z1[0..15]) by repeated squaring of the one-step zero-input map."""
    A = np.zeros((32, 32))
    for k in range(32):
        z0 = np.zeros((16, 1)); z1 = np.zeros((16, 1))
        (z0 if k < 16 else z1)[k % 16, 0] = 1.0
        run_cascade(op, np.zeros((1, 1)), z0, z1)
        A[:16, k] = z0[:, 0]
        A[16:, k] = z1[:, 0]
    return np.linalg.matrix_power(A.astype(np.longdouble), n).astype(np.float64) if n < 64 else _matpow(A, n)


def _matpow(A, n):
    R = np.eye(A.shape[0], dtype=np.longdouble)
    B = A.astype(np.longdouble)
    while n:
        if n & 1:
            R = R @ B
        B = B @ B
        n >>= 1
    return R.astype(np.float64)


def stage_plan(op, tol, hop, bp_margin=1e-6, notch_margin=1e-2):
    """Rows after which sections [0, s) may be dropped from the zero-input fix-up, for s in (8, 10, 12, 14), and the total."""
    c2 = op["c2"]
    rad = np.sqrt(np.abs(c2))            # complex pole pairs: |p|^2 = a2
    rows = np.empty(16)
    for s in range(16):
        t = tol * (bp_margin if s < 8 else notch_margin)
        rows[s] = np.ceil(np.log(t) / np.log(rad[s]))
    cum = np.maximum.accumulate(rows)
    up = lambda v: int(-(-int(v) // hop) * hop)
    return [(up(cum[7]), 8), (up(cum[9]), 10), (up(cum[11]), 12), (up(cum[13]), 14), (up(cum[15]), 16)]


def chunked(op, x, Lc, stages, L, hop, dtype=np.float64):
    """Zero-state chunks + carry + staged fix-up; returns the corrected filtered signal y (gain-free units)."""
    T, C = x.shape
    nch = -(-T // Lc)
    y = np.empty((T, C))
    ends = []
    for j in range(nch):
        lo, hi = j * Lc, min(T, (j + 1) * Lc)
        z0 = np.zeros((16, C), dtype); z1 = np.zeros((16, C), dtype)
        if j == 0:
            z0[:] = op["f0"][:, None].astype(dtype); z1[:] = op["f1"][:, None].astype(dtype)
        y[lo:hi] = run_cascade(op, x[lo:hi], z0, z1, dtype=dtype)
        ends.append(np.concatenate([z0, z1]).astype(np.float64))
    M = transition_matrix(op, Lc)
    s0 = None
    K = stages[-1][0]
    for j in range(1, nch):
        s0 = ends[j - 1] if j == 1 else ends[j - 1] + M @ s0
        lo, hi = j * Lc, min(T, (j + 1) * Lc)
        z0 = s0[:16].copy(); z1 = s0[16:].copy()
        t = 0
        first = 0
        for rows_end, nxt in stages:
            n = min(rows_end, hi - lo) - t
            if n > 0 and first < 16:
                y[lo + t: lo + t + n] += run_cascade(op, np.zeros((n, C)), z0, z1, first=first)
                t += n
            first = nxt
    return y, K


def main():
    fs, C, T = 1000, 6, 42000
    hop, L = fs // 100, fs // 20
    rng = np.random.default_rng(5)
    x = O.synthetic_ecog(T, C, fs, "ecog")[:, :C] + 5000.0 * rng.standard_normal(C)[None, :]   # large per-channel DC offset
    op = lane_operands(fs)
    ref = O.HighGammaExtractor(fs, C).extract_features(x)
    # 1. the kernel's section forms + fresh state == the reference (single chunk)
    z0 = np.repeat(op["f0"][:, None], C, 1); z1 = np.repeat(op["f1"][:, None], C, 1)
    y = run_cascade(op, x, z0, z1)
    f = features_from_y(y, op, L, hop)
    print("single chunk, form56 + fresh state : max|feature - oracle| = %.3e" % np.abs(f - ref).max())
    for tol, Lc in ((1e-11, 24320), (1e-11, 3000), (1e-6, 3000)):
        st = stage_plan(op, tol, hop)
        yc, K = chunked(op, x, Lc, st, L, hop)
        fc = features_from_y(yc, op, L, hop)
        print("chunks of %5d rows, tol %.0e, stages %s : max|feature - oracle| = %.3e" % (Lc, tol, st, np.abs(fc - ref).max()))


if __name__ == "__main__":
    main()
