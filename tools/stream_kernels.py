"""A few streamed packets for an ncu launch list (kernel durations of the latency path)."""
import sys
import numpy as np
sys.path.insert(0, ".")
from delayed_speech_synthesis_b200.engine import HgaEngine
x = np.random.default_rng(1).standard_normal((400, 128))
eng = HgaEngine(1000, 128, precision="exact", device=0)
for i in range(40):
    eng.push(x[i * 10:(i + 1) * 10], np.zeros(128), np.ones(128))
