"""Small driver for ncu captures of the auxiliary kernels added late in round 1: one call each of the rolling log-power
kernel and the raw-ingest channel map on K2-sized data (run under `ncu -k regex:...`)."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from delayed_speech_synthesis_b200 import ops, transforms as TR

T, C = 3_600_000, 128
x = torch.randn((T, C), dtype=torch.float64, device="cuda")
for _ in range(2):
    ops.log_power(x, 1000, 0.05, 0.01)
ident = TR.ChannelMap(TR.ChannelMapSpec.identity(C))
raw = (x * 1000).to(torch.int16)
o = torch.empty((T, C), dtype=torch.float64, device="cuda")
for _ in range(2):
    ident.apply_raw(raw, np.linspace(0.02, 0.3, C), out=o)
torch.cuda.synchronize()
print("ok")
