"""Generate tests/golden/*.npz by EXECUTING the reference in this container (needs /root/reference; run once, commit
the outputs).  Nothing here is imported by the product.

  python tests/golden/make_golden.py

Pieces executed:
  * the reference's compiled extension (oracle/_ref, built from /root/reference/extensions/hga/hga_optimized.pyx):
    compute_log_power_features, and WarmStartFrameBuffer in a keep-alive subprocess (under Cython 3.3 the class aborts
    when a returned slice is released, SURVEY.md §8c — so every returned object is kept and the child leaves via
    os._exit);
  * scipy.signal.sosfilt / sosfilt_zi / iirfilter (the third-party code the reference calls, local/units.py:124-152);
  * the composition of local/units.py:145-161 written out literally with those pieces;
  * the reference's channel-transform classes (local/common.py), cut out of its AST and executed unmodified
    (`python tests/golden/make_golden.py transforms` regenerates only transforms.npz);
  * the reference's packet code (development_amplifier.py BCI2000Package, local/units.py ZMQConnector.interpret_bytes), cut
    out of the AST and executed unmodified (`... make_golden.py packets` regenerates only packets.npz);
  * the reference's Experiment classes (local/common.py:676-880: trial / stimulus indices from StimulusCode), cut out of the
    AST and executed unmodified on synthetic state vectors (`... make_golden.py trials` regenerates only trials.npz).
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parents[1]
sys.path.insert(0, str(ROOT))

CHILD = r"""
import os, sys, numpy as np
sys.path.insert(0, {root!r})
from oracle import build_ref
ref = build_ref.load()
rng = np.random.Generator(np.random.Philox(key=[7, 0]))
chunks = {chunks!r}
C = {C}
fb = ref.WarmStartFrameBuffer(0.05, 0.01, 1000, C)
keep, outs, ins = [], [], []
for n in chunks:
    d = rng.standard_normal((n, C))
    r = fb.insert(d)
    keep.append(r); keep.append(d)
    ins.append(d); outs.append(np.array(r, dtype=np.float64, copy=True))
np.savez({out!r}, chunks=np.array(chunks), **{{f"in{{i}}": a for i, a in enumerate(ins)}}, **{{f"out{{i}}": a for i, a in enumerate(outs)}})
sys.stdout.flush()
os._exit(0)
"""


def framebuffer_golden():
    for name, chunks, C in (("framebuffer_a", [10, 10, 10, 20, 40, 40, 70], 3), ("framebuffer_b", [60, 10, 25, 13], 2),
                            ("framebuffer_c", [7, 9, 33, 10], 2)):
        out = str(HERE / f"{name}.npz")
        code = CHILD.format(root=str(ROOT), chunks=chunks, C=C, out=out)
        subprocess.run([sys.executable, "-c", code], check=True)


def reference_transform_classes():
    """The reference's own channel-transform classes, executed from where they lie: /root/reference/local/common.py cannot
    be imported here (h5py, sklearn imports at its top), so the five class definitions are cut out of its AST and
    executed unmodified in a namespace that holds what they use (numpy, scipy.ndimage.binary_dilation, typing.List)."""
    import ast
    from typing import List
    from scipy.ndimage import binary_dilation
    src = Path("/root/reference/local/common.py").read_text()
    want = ["SelectElectrodesFromBothGrids", "SelectElectrodesOverSpeechAreas", "BadChannelCorrection", "CommonAverageReferencing",
            "ZScoreNormalization"]
    ns = {"np": np, "List": List, "binary_dilation": binary_dilation}
    for node in ast.parse(src).body:
        if isinstance(node, ast.ClassDef) and node.name in want:
            exec(compile(ast.Module(body=[node], type_ignores=[]), "/root/reference/local/common.py", "exec"), ns)
    return {k: ns[k] for k in want}


def transforms_golden():
    """prepare_corpus.py:147-175 and decode_online.py:64-82 wiring, on seeded data (tests/golden/transforms.npz)."""
    R = reference_transform_classes()
    rng = np.random.Generator(np.random.Philox(key=[20241104, 7]))
    x = rng.standard_normal((97, 128)) * 30 + 20 * np.sin(2 * np.pi * 60 * np.arange(97) / 1000)[:, None]
    speech_grid = np.flip(np.arange(64, dtype=np.int16).reshape((8, 8)) + 1, axis=0)
    motor_grid = np.flip(np.arange(64, dtype=np.int16).reshape((8, 8)) + 65, axis=0)
    layout = np.arange(128) + 1
    both, speech = R["SelectElectrodesFromBothGrids"](), R["SelectElectrodesOverSpeechAreas"]()
    car = R["CommonAverageReferencing"](exclude_channels=[19, 38, 48, 52], grids=[speech_grid, motor_grid], layout=layout)
    bcc = R["BadChannelCorrection"](bad_channels=[19, 38, 48, 52, 1, 64, 100], grids=[speech_grid, motor_grid], layout=layout)
    a = both(x)
    b = car(a)
    c = speech(b)
    f = rng.standard_normal((61, 128)) * 0.4 + 3.0            # stands for (F, C) features
    out = {"x": x, "both_mapping": np.asarray(both.grid_mapping), "speech_mapping": np.asarray(speech.speech_grid_mapping),
           "after_both": a, "after_car": b, "after_speech": c, "exclude": np.array([19, 38, 48, 52]),
           "car_appl": np.array(car.selection_masks_application), "car_comp": np.array(car.selection_masks_computation),
           "speech_grid": speech_grid, "motor_grid": motor_grid, "layout": layout,
           "bad": np.array([19, 38, 48, 52, 1, 64, 100]), "f": f, "after_bcc": bcc(f),
           "bcc_loc": np.array([int(p[0][0]) for p in bcc.patches]),
           "bcc_neigh": np.array([np.pad(p[1], (0, 8 - len(p[1])), constant_values=-1) for p in bcc.patches])}
    np.savez(HERE / "transforms.npz", **out)


def packets_golden():
    """The reference's own packet code, cut out of its AST and executed unmodified: BCI2000Package (development_amplifier.py
    :14-25) packs seeded payloads, ZMQConnector.interpret_bytes (local/units.py:78-82) decodes them (tests/golden/packets.npz)."""
    import ast
    import struct
    import types
    ns = {"struct": struct, "np": np}
    for node in ast.parse(Path("/root/reference/development_amplifier.py").read_text()).body:
        if isinstance(node, ast.ClassDef) and node.name == "BCI2000Package":
            exec(compile(ast.Module(body=[node], type_ignores=[]), "/root/reference/development_amplifier.py", "exec"), ns)
    for node in ast.parse(Path("/root/reference/local/units.py").read_text()).body:
        if isinstance(node, ast.ClassDef) and node.name == "ZMQConnector":
            fn = next(n for n in node.body if isinstance(n, ast.FunctionDef) and n.name == "interpret_bytes")
            exec(compile(ast.Module(body=[fn], type_ignores=[]), "/root/reference/local/units.py", "exec"), ns)
    fake_self = types.SimpleNamespace(STATE=types.SimpleNamespace(header=struct.Struct('=BBB HH')))
    rng = np.random.Generator(np.random.Philox(key=[20241104, 11]))
    out = {}
    for name, (c, n) in {"a": (128, 10), "b": (5, 3), "c": (64, 40), "d": (1, 1)}.items():
        payload = (rng.standard_normal((c, n)) * 30).astype(np.float32)
        pkt = ns["BCI2000Package"](c, n).pack(payload)
        out[f"packet_{name}"] = np.frombuffer(pkt, dtype=np.uint8)
        out[f"decoded_{name}"] = ns["interpret_bytes"](fake_self, pkt)
    np.savez(HERE / "packets.npz", **out)


def trial_cases():
    """Synthetic BCI2000 `StimulusCode` state vectors (the committed inputs of trials.npz): regular runs of both paradigms,
    a recording that ends inside a stimulus, back-to-back non-zero codes, a single trial, leading / trailing rest."""
    rng = np.random.Generator(np.random.Philox(key=[20241104, 5]))
    cases = {}

    def paradigm(n_trials, n_codes, lead, stim, speak, rest, jitter, tail, open_end=False, back_to_back=False):
        parts, codes = [np.zeros(lead, np.int16)], rng.integers(1, n_codes + 1, n_trials)
        for i, cde in enumerate(codes):
            parts.append(np.full(stim + int(rng.integers(0, jitter + 1)), cde, np.int16))
            if back_to_back and i % 3 == 1:
                parts.append(np.full(stim // 2, (cde % n_codes) + 1, np.int16))
            parts.append(np.zeros(speak + int(rng.integers(0, jitter + 1)) + rest, np.int16))
        if open_end:
            parts.append(np.full(stim // 2, codes[0], np.int16))
        else:
            parts.append(np.zeros(tail, np.int16))
        return np.concatenate(parts)

    cases["syll_regular"] = ("SyllableRepetition", paradigm(24, 12, 3000, 700, 1800, 0, 0, 4000), 12)
    cases["syll_jitter"] = ("SyllableRepetition", paradigm(40, 12, 1234, 650, 1500, 0, 300, 2500), 12)
    cases["syll_two_trials"] = ("SyllableRepetition", paradigm(2, 4, 100, 50, 120, 0, 0, 500), 4)
    cases["syll_short_tail"] = ("SyllableRepetition", paradigm(6, 6, 500, 400, 900, 0, 50, 10), 6)
    cases["key_regular"] = ("KeywordReading", paradigm(30, 6, 2000, 2000, 1000, 500, 0, 3000), 6)
    cases["key_jitter"] = ("KeywordReading", paradigm(50, 6, 10, 1500, 700, 200, 400, 1), 6)
    cases["key_open_end"] = ("KeywordReading", paradigm(12, 6, 800, 1000, 600, 0, 100, 0, open_end=True), 6)
    cases["key_back_to_back"] = ("KeywordSynthesis", paradigm(15, 6, 400, 900, 500, 0, 100, 700, back_to_back=True), 6)
    cases["key_single"] = ("KeywordReading", paradigm(1, 6, 300, 500, 400, 0, 0, 200), 6)
    cases["key_none"] = ("KeywordReading", np.zeros(5000, np.int16), 6)
    return cases


def trials_golden():
    """The reference's Experiment classes (local/common.py:676-880), cut out of the AST and executed unmodified on the
    synthetic state vectors above; BCI2000MatFile.trial_indices' min_trial_length clamp (:432-445) written out literally."""
    import ast
    from abc import ABC, abstractmethod
    from functools import partial
    from typing import Callable, Dict, List, Optional, Tuple, Union
    src = Path("/root/reference/local/common.py").read_text()
    want = ["Experiment", "SyllableRepetition", "KeywordReading", "ExperimentMapping"]
    ns = {"np": np, "os": os, "ABC": ABC, "abstractmethod": abstractmethod, "Dict": Dict, "List": List, "Optional": Optional, "Tuple": Tuple,
          "Union": Union, "Callable": Callable, "partial": partial}
    for node in ast.parse(src).body:
        if isinstance(node, ast.ClassDef) and node.name in want:
            exec(compile(ast.Module(body=[node], type_ignores=[]), "/root/reference/local/common.py", "exec"), ns)
    out = {}
    for name, (paradigm, code, n_codes) in trial_cases().items():
        stimuli = [f"stim{j + 1:02d}" for j in range(n_codes)]
        cls = ns["ExperimentMapping"].get_experiment_class(f"/data/day1/{paradigm}_Run01.mat")
        exp = cls(code, stimuli)
        out[f"{name}.paradigm"] = np.array(paradigm)
        out[f"{name}.code"] = code
        out[f"{name}.n_codes"] = np.array(n_codes)
        for kind, fn in (("trials", exp.get_trial_indices), ("stimuli", exp.get_stimuli_indices)):
            try:
                res = fn()
                out[f"{name}.{kind}.labels"] = np.array([r[0] for r in res], dtype="U16")
                out[f"{name}.{kind}.bounds"] = np.array([[int(r[1]), int(r[2])] for r in res], dtype=np.int64).reshape(-1, 2)
            except Exception as exc:   # the reference raises on degenerate vectors (e.g. IndexError without any trial)
                out[f"{name}.{kind}.raises"] = np.array(type(exc).__name__)
        try:
            bl = exp.get_webfm_baseline_windows(1000)
            out[f"{name}.baseline.bounds"] = np.array([[int(r[1]), int(r[2])] for r in bl], dtype=np.int64).reshape(-1, 2)
            run = exp.get_experiment_run_indices()
            out[f"{name}.run"] = np.array([int(run[1]), int(run[2])], dtype=np.int64)
        except Exception as exc:
            out[f"{name}.run.raises"] = np.array(type(exc).__name__)
    np.savez_compressed(HERE / "trials.npz", **out)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "trials":
        trials_golden()
        print("golden written: trials.npz")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "packets":
        packets_golden()
        print("golden written: packets.npz")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "transforms":
        transforms_golden()
        print("golden written: transforms.npz")
        return
    from scipy.signal import iirfilter, sosfilt, sosfilt_zi
    from oracle import build_ref

    ref = build_ref.load()
    assert ref is not None, "reference extension unavailable (needs /root/reference)"
    rng = np.random.Generator(np.random.Philox(key=[20241104, 99]))

    # 1. compute_log_power_features of the compiled reference
    lp = {}
    for i, (rows, C, sr, wl, ws) in enumerate([(1000, 4, 1000, 0.05, 0.01), (50, 2, 1000, 0.05, 0.01), (257, 5, 1000, 0.05, 0.01),
                                               (600, 3, 2000, 0.05, 0.01), (700, 2, 1000, 0.05, 0.0125), (333, 2, 1000, 0.032, 0.008)]):
        d = rng.standard_normal((rows, C)) * (1.0 + i)
        lp[f"in{i}"], lp[f"par{i}"] = d, np.array([sr, wl, ws])
        lp[f"out{i}"] = np.array(ref.compute_log_power_features(d, sr, wl, ws))
    np.savez(HERE / "log_power.npz", **lp)

    # 2. filters: design, zi, sosfilt outputs and final states (scipy as installed here: see versions below)
    import scipy
    filt = {"scipy_version": np.array(scipy.__version__)}
    for fs in (1000, 2000):
        nyq = fs / 2
        bp = iirfilter(8, np.array([70, 170]) / nyq, btype="bandpass", ftype="butter", output="sos")
        nt = iirfilter(8, np.array([118, 122]) / nyq, btype="bandstop", ftype="butter", output="sos")
        x = rng.standard_normal((400, 3))
        zb = np.repeat(sosfilt_zi(bp), 3, axis=-1).reshape(8, 2, 3)
        zn = np.repeat(sosfilt_zi(nt), 3, axis=-1).reshape(8, 2, 3)
        y1, s1 = sosfilt(bp, x, axis=0, zi=zb)
        y2, s2 = sosfilt(nt, y1, axis=0, zi=zn)
        filt.update({f"bp{fs}": bp, f"nt{fs}": nt, f"zib{fs}": sosfilt_zi(bp), f"zin{fs}": sosfilt_zi(nt), f"x{fs}": x,
                     f"y1_{fs}": y1, f"s1_{fs}": s1, f"y2_{fs}": y2, f"s2_{fs}": s2})
    np.savez(HERE / "filters.npz", **filt)

    # 3. the whole path, local/units.py:145-161 literally: offline single call and 10-sample packets, + z-score
    fs, C, T = 1000, 6, 1200
    x = rng.standard_normal((T, C)) * 30 + 20 * np.sin(2 * np.pi * 120 * np.arange(T) / fs)[:, None]
    bp, nt = filt["bp1000"], filt["nt1000"]
    zb = np.repeat(sosfilt_zi(bp), C, axis=-1).reshape(8, 2, C)
    zn = np.repeat(sosfilt_zi(nt), C, axis=-1).reshape(8, 2, C)
    d, _ = sosfilt(bp, x, axis=0, zi=zb)
    d, _ = sosfilt(nt, d, axis=0, zi=zn)                       # frame buffer case 1: data passes through unchanged
    off = np.array(ref.compute_log_power_features(d, fs, 0.05, 0.01))
    mu, sd = np.mean(off, axis=0), np.std(off, axis=0)         # prepare_corpus.py:198-199
    # streaming: case 2 then case 3 (restated buffer; framebuffer_*.npz pins the restatement against the compiled class)
    hs, fh, rem, first, frames = zb.copy(), zn.copy(), np.zeros((40, C)), True, []
    for t in range(0, 300, 10):
        c, hs = sosfilt(bp, x[t:t + 10], axis=0, zi=hs)
        c, fh = sosfilt(nt, c, axis=0, zi=fh)
        buf = np.concatenate([np.zeros((40, C)), c]) if first else np.concatenate([rem, c])
        first, rem = False, buf[-40:]
        frames.append(np.array(ref.compute_log_power_features(np.ascontiguousarray(buf), fs, 0.05, 0.01)))
    np.savez(HERE / "pipeline.npz", x=x, offline=off, mu=mu, sigma=sd, offline_z=(off - mu) / sd,
             stream=np.concatenate(frames), stream_hg_state=hs, stream_fh_state=fh)
    framebuffer_golden()
    transforms_golden()
    packets_golden()
    print("golden written:", sorted(p.name for p in HERE.glob("*.npz")))


if __name__ == "__main__":
    main()
