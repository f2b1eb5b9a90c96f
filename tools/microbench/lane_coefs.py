"""Writes the filter tables the lane_bench harness reads (tools/microbench/lane_bench.cu): for fs in (1000, 2000) the
band-pass / notch SOS matrices and their sosfilt_zi, float64, as lane_coefs_<fs>.bin = [sos_bp 8x6 | zi_bp 8x2 | sos_notch 8x6 | zi_notch 8x2]."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from delayed_speech_synthesis_b200 import filters  # noqa: E402

here = Path(__file__).resolve().parent
for fs in (1000, 2000):
    hg, fh = filters.design_hga_filters(fs)
    blob = np.concatenate([hg.ravel(), filters.initial_state(hg).ravel(), fh.ravel(), filters.initial_state(fh).ravel()]).astype(np.float64)
    assert blob.size == 128
    (here / f"lane_coefs_{fs}.bin").write_bytes(blob.tobytes())
print("ok")
