"""Target of the ncu capture of the fused chain + extraction kernel (profiles/r02_ncu_chain_f32.csv): K2 through the
reference's pre-transform chain (129 raw columns -> 64 channels), one warm-up call and one captured call."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from delayed_speech_synthesis_b200 import transforms as TR
from delayed_speech_synthesis_b200.engine import HgaEngine

fs, c_raw, C, T = 1000, 129, 64, 3_600_000
x = torch.randn((T, c_raw), dtype=torch.float64, device="cuda")
layout = np.arange(128) + 1
grids = [np.flip(layout[:64].reshape(8, 8), axis=0), layout[64:].reshape(8, 8)]
pre = [TR.ChannelSelection(np.arange(128)), TR.CommonAverageReferencing([19, 38, 77], grids, layout), TR.ChannelSelection(np.arange(64, 128))]
cmap = TR.ChannelMap(TR.fold(pre, c_raw), 0)
eng = HgaEngine(fs, C, precision="f32", device=0)
for _ in range(3):
    out = eng.extract_offline_mapped(x, cmap)
torch.cuda.synchronize()
print("fused", eng.last_fused, eng.last_plan())


def timed(fn, n=10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


if len(sys.argv) > 1 and sys.argv[1] == "time":
    xm = torch.empty((T, C), dtype=torch.float64, device="cuda")
    out2 = torch.empty_like(out)
    print("fused ms", timed(lambda: eng.extract_offline_mapped(x, cmap, out=out)))
    print("two-pass ms", timed(lambda: eng.extract_offline(cmap.apply(x, out=xm), out=out2)))
    eng.set_timing(True)
    eng.extract_offline_mapped(x, cmap, out=out)
    print("main kernel / rest ms", eng.last_kernel_split_ms())
    print("max |fused - two-pass|", float((out - out2).abs().max()))
