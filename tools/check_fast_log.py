"""NumPy emulation of fast_log (csrc/hga_kernels_pipe.cuh): series degree / ln2 split check against numpy.log."""
import numpy as np

rng = np.random.default_rng(0)
x = np.concatenate([10 ** rng.uniform(-2, 15, 2_000_000), [0.01, 1.0, 2.0, np.sqrt(2), 1e15]])
m, e = np.frexp(x)
m, e = m * 2, e - 1
big = m > 1.4142135623730951
m, e = np.where(big, m * 0.5, m), np.where(big, e + 1, e).astype(float)
s = (m - 1) / (m + 1)
z = s * s
q = np.zeros_like(z)
for k in range(11, 0, -1):
    q = z * (1.0 / (2 * k + 1) + q)
res = (e * float.fromhex("0x1.62e42fee00000p-1") + 2 * s) + (2 * s * q + e * float.fromhex("0x1.a39ef35793c76p-33"))
err = np.abs(res - np.log(x))
print("max abs", err.max(), "max ulp", (err / np.spacing(np.abs(np.log(x)))).max())
