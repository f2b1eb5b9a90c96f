"""A/B of the baseline flow on one K2 recording (128 ch x 1 h): extraction with the statistics left behind by the kernels
(hga_extract_offline_stats) against extraction followed by the statistics pass (hga_stats_accumulate)."""
import sys

import torch

sys.path.insert(0, ".")
from delayed_speech_synthesis_b200 import ops
from delayed_speech_synthesis_b200.engine import HgaEngine

fs, C, T = 1000, 128, 3_600_000
x = torch.randn((T, C), dtype=torch.float64, device="cuda")


def timed(fn, n=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for prec in ("f32", "f64"):
    eng = HgaEngine(fs, C, precision=prec, device=0)
    out = torch.empty((eng.num_frames(T), C), dtype=torch.float64, device="cuda")
    st = ops.ChannelStats(C, 0)
    a = timed(lambda: eng.extract_offline_stats(x, st, out=out))
    b = timed(lambda: (eng.extract_offline(x, out=out), st.update(out)))
    c = timed(lambda: eng.extract_offline(x, out=out))
    print(f"{prec}: fused {a:.4f} ms, extraction + statistics pass {b:.4f} ms, extraction alone {c:.4f} ms")
