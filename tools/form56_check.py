"""Design check for the 56-slot cascade form of k_hga_pipe2: the band-pass numerator (1+z^-1)^8 (1-z^-1)^8 is
re-paired into eight [1, 0, -1] sections (3 FP slots each instead of 4), sections are interleaved [A, A, U, U] per filter
warp, the cascade starts from ZERO state and the zero-input response of the reference's sosfilt_zi state is added to the
output.  Prints max-abs feature error vs the float64 oracle for float64 and emulated float32 arithmetic."""
import sys
import numpy as np
sys.path.insert(0, ".")
from oracle import hga_oracle as O
from delayed_speech_synthesis_b200 import filters


def run(dtype, x, secs, zir):
    x = x.astype(dtype)
    C = x.shape[1]
    st = [(np.zeros(C, dtype), np.zeros(C, dtype)) for _ in secs]
    y = np.empty_like(x)
    cs = [(k, dtype(a1), dtype(a2), dtype(b1)) for k, a1, a2, b1 in secs]
    for t in range(len(x)):
        v = x[t]
        for i, (k, a1, a2, b1) in enumerate(cs):
            p, w = st[i]
            if k == "A":      # y = x - v; v' = a1 y + w; w' = a2 y + x     (negated DF2T state, numerator [1, 0, -1])
                o = v - p
                st[i] = (a1 * o + w, a2 * o + v)
            else:             # numerator [1, b1, 1]
                o = v + p
                st[i] = ((b1 * v + w) + (-a1) * o, (-a2) * o + v)
            v = o
        y[t] = v
    y = y.astype(np.float64)
    y[:len(zir)] += zir[:len(y), None]
    return y


def main():
    fs, C, T = 1000, 8, 6000
    x = O.synthetic_ecog(T, C, fs, "ecog")
    bp, nt = filters.create_filter_sos(fs, 70, 170, 8), filters.create_filter_sos(fs, 122, 118, 8)
    ref = O.HighGammaExtractor(fs, C)
    want = ref.extract_features(x)
    G = bp[0, 0] * nt[0, 0]
    from scipy.signal import sosfilt, sosfilt_zi
    z = np.zeros(T)
    r1, _ = sosfilt(bp, z, zi=sosfilt_zi(bp))
    # zero-input response of the whole chain: band-pass ZIR through the notch (its own zi) 
    zir, _ = sosfilt(nt, r1, zi=sosfilt_zi(nt))
    A = [("A", s[4], s[5], 0.0) for s in bp]
    U = [("U", s[4], s[5], s[1] / s[0]) for s in nt]
    for name, order in (("A8 U8", A + U), ("[AAUU]x4", sum([A[2 * g:2 * g + 2] + U[2 * g:2 * g + 2] for g in range(4)], []))):
        for dt in (np.float64, np.float32):
            y = run(dt, x, order, zir / G) * G
            f = O.compute_log_power_features(y, fs, 0.05, 0.01)
            print(f"{name:10s} {dt.__name__:8s} max-abs feature error {np.abs(f - want).max():.3e}")


if __name__ == "__main__":
    main()
