"""End-to-end (host buffers in, host features out) of one K2 recording fed as raw counts of each element type through
hga_extract_offline_raw_host, beside the float64 call.  Pinned host memory, wall clock around the call, median of 5."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from delayed_speech_synthesis_b200 import transforms as TR
from delayed_speech_synthesis_b200.engine import HgaEngine

T, C, fs = 3_600_000, 128, 1000
eng = HgaEngine(fs, C)
F = eng.num_frames(T)
ident = TR.ChannelMap(TR.ChannelMapSpec.identity(C), eng.device)
gain = np.linspace(0.05, 0.15, C)
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn((T, C), dtype=torch.float64, device="cuda", generator=g) * 30.0
oh = torch.empty((F, C), dtype=torch.float64, pin_memory=True)
on = oh.numpy()
mu, sg = np.zeros(C), np.ones(C)
res = {}


def timed(fn, n=5):
    fn()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts))


xh = torch.empty((T, C), dtype=torch.float64, pin_memory=True)
xh.copy_(x)
xn = xh.numpy()
dt = timed(lambda: eng.extract_offline(xn, None, mu, sg, out=on))
res["float64 (hga_extract_offline_host)"] = {"ms": dt * 1e3, "G ch-samples/s": C * T / dt / 1e9, "h2d_bytes": T * C * 8}
del xh, xn
for tdt in (torch.float32, torch.int32, torch.int16):
    counts = torch.clamp(torch.round(x / torch.from_numpy(gain).cuda()), -32000, 32000).to(tdt)
    ch = torch.empty((T, C), dtype=tdt, pin_memory=True)
    ch.copy_(counts)
    cn = ch.numpy()
    dt = timed(lambda: eng.extract_offline_raw(cn, ident, gain, None, mu, sg, out=on))
    want = eng.extract_offline(ident.apply_raw(counts, gain), None, torch.from_numpy(mu).cuda(), torch.from_numpy(sg).cuda())
    res[f"{str(tdt).split('.')[-1]} counts x gain (hga_extract_offline_raw_host)"] = {
        "ms": dt * 1e3, "G ch-samples/s": C * T / dt / 1e9, "h2d_bytes": T * C * ch.element_size(), "bit_identical_to_device_path": bool(torch.equal(oh.cuda(), want))}
    del ch, cn, counts, want
print(json.dumps({"workload": "K2 128 ch x 1 h @ 1 kHz", "d2h_bytes": F * C * 8, "calls": res}, indent=1))
