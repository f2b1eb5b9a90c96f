"""NumPy emulation of the device logs (csrc/hga_async.cuh tab_log, and the atanh-series fast_log it replaced):
series degree / table / ln2 split check against numpy.log (long double)."""
import numpy as np

rng = np.random.default_rng(0)
x = np.concatenate([10 ** rng.uniform(-2, 15, 2_000_000), rng.uniform(0.01, 4, 1_000_000), [0.01, 1.0, 2.0, np.sqrt(2), 1e15]])
ref = np.log(x.astype(np.longdouble))
LN2HI, LN2LO = float.fromhex("0x1.62e42fee00000p-1"), float.fromhex("0x1.a39ef35793c76p-33")

# fast_log: atanh series
m, e = np.frexp(x)
m, e = m * 2, e - 1
big = m > 1.4142135623730951
m, e = np.where(big, m * 0.5, m), np.where(big, e + 1, e).astype(float)
s = (m - 1) / (m + 1)
z = s * s
q = np.zeros_like(z)
for k in range(11, 0, -1):
    q = z * (1.0 / (2 * k + 1) + q)
res = (e * LN2HI + 2 * s) + (2 * s * q + e * LN2LO)
err = np.abs(res - ref).astype(float)
print("fast_log: max abs", err.max(), "max ulp", (err / np.spacing(np.abs(np.log(x)))).max())

# tab_log: 128-entry table, degree-6 log1p series (plain float64 products stand in for the device FMAs)
N = 128
i = np.arange(N, dtype=np.longdouble)
rc = (1 / (1 + (i + 0.5) / N)).astype(np.float64)
nl = (-np.log(rc.astype(np.longdouble))).astype(np.float64)
m, e = np.frexp(x)
m, e = m * 2, (e - 1).astype(float)
idx = np.minimum(((m - 1) * N).astype(int), N - 1)
r = ((m.astype(np.longdouble) * rc[idx].astype(np.longdouble)) - 1).astype(np.float64)   # fma: one rounding
qv = r * (-1 / 6) + 1 / 5
qv = r * qv - 1 / 4
qv = r * qv + 1 / 3
qv = r * qv - 0.5
lo = (r * r) * qv + r
res = (e * LN2HI + nl[idx]) + (e * LN2LO + lo)
err = np.abs(res - ref).astype(float)
print("tab_log : max abs", err.max(), "max ulp", (err / np.spacing(np.abs(np.log(x)))).max())
small = x < 4
print("tab_log on [0.01, 4): max abs", err[small].max())
