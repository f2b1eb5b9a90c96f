"""Target of the compute-sanitizer runs (profiles/r02_sanitizer_*.log): every kernel family once, at sizes the tools finish in
minutes — a K1-size recording through the chunked lane kernels (float64, packed float32, mixed; several carried chunks so the
carry and head fix-up kernels run), ragged segments through the generic float64 kernel, the host-buffer call, the statistics
kernels and 100 streamed packets on one handle.

    compute-sanitizer --tool memcheck  python tools/sanitize_target.py
    compute-sanitizer --tool racecheck python tools/sanitize_target.py
"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from delayed_speech_synthesis_b200.engine import HgaEngine

fs, C, secs = 1000, 128, 60
T = fs * secs
g = torch.Generator(device="cuda").manual_seed(7)
x = torch.randn((T, C), dtype=torch.float64, device="cuda", generator=g)
mu = torch.zeros(C, dtype=torch.float64, device="cuda")
sigma = torch.ones(C, dtype=torch.float64, device="cuda")
ref = None
for prec in ("f64", "f32", "mixed", "exact"):
    eng = HgaEngine(fs, C, precision=prec, device=0)
    out = eng.extract_offline(x, None, mu, sigma)
    torch.cuda.synchronize()
    if ref is None:
        ref = out
    print(prec, "plan", eng.last_plan(), "max |diff| vs f64", float((out - ref).abs().max()))
    # ragged fresh segments (shorter than the carry: whole-segment items), then overlapping ones
    segs = [(0, 777), (1000, 1050), (5000, 21000), (30000, 30050), (40000, 59990)]
    o2 = eng.extract_offline(x, segs, mu, sigma)
    # host-buffer call on a short slice (pageable numpy in / out)
    xh = x[:20000].cpu().numpy()
    o3 = eng.extract_offline(xh, None, mu.cpu().numpy(), sigma.cpu().numpy())
    assert np.isfinite(o3).all() and torch.isfinite(o2).all()
    del eng

# baseline statistics: accumulate, merge, finalize, then the z-score kernel
from delayed_speech_synthesis_b200 import ops
st = ops.ChannelStats(C, 0).update(ref[:3000]).update(ref[3000:])
st2 = ops.ChannelStats(C, 0).update(ref[:100])
st.merge(st2)
m, s_ = st.finalize(as_numpy=False)
z = ops.zscore(ref, m, s_)
assert torch.isfinite(z).all()

# streaming: 100 packets of 10 samples on one handle (device push), then 100 host pushes
eng = HgaEngine(fs, C, precision="exact", device=0)
for i in range(100):
    eng.push(x[i * 10:(i + 1) * 10], mu, sigma)
xs = x[:2000].cpu().numpy()
mh, sh = mu.cpu().numpy(), sigma.cpu().numpy()
for i in range(100):
    eng.push(xs[i * 10:(i + 1) * 10], mh, sh)
torch.cuda.synchronize()
print("ok")
