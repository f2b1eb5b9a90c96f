"""HBM-bound auxiliary kernels of the path against the copy roofline: channel map, baseline statistics,
packet decode.  Prints achieved GB/s (algorithmic bytes / CUDA-event time) and the fraction of MEASURED_PEAKS.json:hbm_gbs."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, ".")
from delayed_speech_synthesis_b200 import transforms as TR
from delayed_speech_synthesis_b200.ops import ChannelStats

ROOT = Path(__file__).resolve().parent.parent
peak = 6553.9
try:
    peak = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
except Exception:
    pass


def timed(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(n):
        fn()
    ev[1].record()
    torch.cuda.synchronize()
    return ev[0].elapsed_time(ev[1]) / n * 1e-3


T, C = 3_600_000, 128
x = torch.randn((T, C), dtype=torch.float64, device="cuda")
layout = np.arange(128) + 1
grids = [np.flip(layout[:64].reshape(8, 8), 0), np.flip(layout[64:].reshape(8, 8), 0)]
rng = np.random.default_rng(1)
chain = [TR.ChannelSelection(rng.permutation(128)), TR.CommonAverageReferencing([19, 38, 48, 52], grids, layout), TR.ChannelSelection(np.arange(64))]
out = {}
for name, ch, c_out in (("select+CAR+select 128->64", chain, 64), ("select+CAR 128->128", chain[:2], 128), ("gather 128->128", chain[:1], 128)):
    m = TR.ChannelMap(TR.fold(ch, C))
    o = torch.empty((T, c_out), dtype=torch.float64, device="cuda")
    s = timed(lambda: m.apply(x, out=o))
    gb = T * 8 * (C + c_out) / s / 1e9
    out[name] = {"ms": s * 1e3, "GB/s": gb, "frac_hbm": gb / peak}
gain = np.linspace(0.02, 0.3, C)
ident = TR.ChannelMap(TR.ChannelMapSpec.identity(C))
o = torch.empty((T, C), dtype=torch.float64, device="cuda")
for dt, nb in ((torch.int16, 2), (torch.float32, 4)):
    raw = (x * 1000).to(dt)
    s = timed(lambda: ident.apply_raw(raw, gain, out=o))
    gb = T * C * (nb + 8) / s / 1e9
    out[f"raw ingest signal*gain {str(dt).split('.')[-1]} -> f64 128->128"] = {"ms": s * 1e3, "GB/s": gb, "frac_hbm": gb / peak}
    del raw
del o
f = torch.randn((360_000 * 8, C), dtype=torch.float64, device="cuda")   # features of 8 one-hour sessions
st = ChannelStats(C, 0)
s = timed(lambda: ChannelStats(C, 0).update(f))
out["baseline stats (n, mean, M2) of 2.88 M x 128 features"] = {"ms": s * 1e3, "GB/s": f.numel() * 8 / s / 1e9, "frac_hbm": f.numel() * 8 / s / 1e9 / peak}
from delayed_speech_synthesis_b200 import ops
mu_t, sg_t = torch.zeros(C, dtype=torch.float64, device="cuda"), torch.ones(C, dtype=torch.float64, device="cuda")
s = timed(lambda: ops.log_power(x, 1000, 0.05, 0.01), n=5, warm=2)      # includes the host-side window-index arithmetic of the call
F = (T - 50) // 10 + 1
gb = (T * C * 8 + F * C * 8) / s / 1e9
out["compute_log_power_features (k_frames_roll, exact row-order sums) 3.6 M x 128"] = {"ms": s * 1e3, "GB/s": gb, "frac_hbm": gb / peak}
s = timed(lambda: ops.zscore(f, mu_t, sg_t))
gb = 2 * f.numel() * 8 / s / 1e9
out["zscore 2.88 M x 128"] = {"ms": s * 1e3, "GB/s": gb, "frac_hbm": gb / peak}
print(json.dumps({"peak_hbm_gbs": peak, "kernels": out}, indent=1))
