"""Where a streamed packet's latency goes: Python wrapper vs C call vs kernel (10 x 128 packets)."""
import ctypes as C
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from delayed_speech_synthesis_b200.engine import HgaEngine

fs, Cn, n, N = 1000, 128, 10, 5000
x = np.random.default_rng(1).standard_normal((N * n, Cn))
mu, sg = np.zeros(Cn), np.ones(Cn)
eng = HgaEngine(fs, Cn, precision="exact", device=0)
lib, h = eng._lib, eng._h
out = np.empty((4, Cn))
nf = C.c_int64(0)


def pct(v):
    v = np.asarray(v[500:]) * 1e6
    return f"p50 {np.percentile(v, 50):6.1f} us  p99 {np.percentile(v, 99):6.1f} us"


t = []
for i in range(N):
    p = x[i * n:(i + 1) * n]
    t0 = time.perf_counter(); eng.push(p, mu, sg); t.append(time.perf_counter() - t0)
print("engine.push (python wrapper + C call)      ", pct(t))
t = []
for i in range(N):
    p = x[i * n:(i + 1) * n]
    t0 = time.perf_counter()
    lib.hga_stream_push_host(h, p.ctypes.data, n, Cn, mu.ctypes.data, sg.ctypes.data, out.ctypes.data, Cn, C.byref(nf))
    t.append(time.perf_counter() - t0)
print("raw ctypes hga_stream_push_host             ", pct(t))
xd = torch.from_numpy(x).cuda(); od = torch.empty((4, Cn), dtype=torch.float64, device="cuda")
mud, sgd = torch.from_numpy(mu).cuda(), torch.from_numpy(sg).cuda()
st = torch.cuda.current_stream().cuda_stream
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
torch.cuda.synchronize()
ks = []
for i in range(2000):
    ev[0].record()
    lib.hga_stream_push(h, xd[i * n:(i + 1) * n].data_ptr(), n, Cn, mud.data_ptr(), sgd.data_ptr(), od.data_ptr(), Cn, C.byref(nf), st)
    ev[1].record(); ev[1].synchronize()
    ks.append(ev[0].elapsed_time(ev[1]) * 1e-3)
print("k_stream_push on the device (CUDA events)   ", pct(ks))
t = []
for i in range(N):
    t0 = time.perf_counter(); torch.cuda.current_stream().synchronize(); t.append(time.perf_counter() - t0)
print("empty stream synchronize                    ", pct(t))
# warm-cache kernel time: 200 device-buffer pushes back to back, no host synchronisation in between
torch.cuda.synchronize()
for rep in range(3):
    ev[0].record()
    for i in range(200):
        lib.hga_stream_push(h, xd[i * n:(i + 1) * n].data_ptr(), n, Cn, mud.data_ptr(), sgd.data_ptr(), od.data_ptr(), Cn, C.byref(nf), st)
    ev[1].record(); ev[1].synchronize()
print(f"back-to-back device pushes: {ev[0].elapsed_time(ev[1]) / 200 * 1e3:.1f} us per packet (kernel + launch throughput)")
