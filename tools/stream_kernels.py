"""A few streamed packets for an ncu launch list (kernel durations of the latency path): host buffers (mapped pinned
memory, zero-copy) first, then device buffers."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from delayed_speech_synthesis_b200.engine import HgaEngine
x = np.random.default_rng(1).standard_normal((800, 128))
eng = HgaEngine(1000, 128, precision="exact", device=0)
for i in range(40):
    eng.push(x[i * 10:(i + 1) * 10], np.zeros(128), np.ones(128))
xd = torch.from_numpy(x).cuda()
mu, sg = torch.zeros(128, dtype=torch.float64, device="cuda"), torch.ones(128, dtype=torch.float64, device="cuda")
for i in range(40, 80):
    eng.push(xd[i * 10:(i + 1) * 10], mu, sg)
torch.cuda.synchronize()
