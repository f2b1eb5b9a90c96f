#!/usr/bin/env python
"""bench.py — HGA channel-samples/s on B200 (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--precision f32|f64|mixed]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

N = 1 (BASELINE config 2, "k2"): a step = one pass of the hot path (IIR cascade + framing + log-power + z-score) over one
resident recording, 128 ch x 1 h @ 1 kHz float64 (3.69 GB > L2: no flush needed between steps).  `value` = device-resident
throughput, `e2e` = the same call on HOST buffers (copies inside the timed region).  Side entries of the same line: the other
precisions, short recordings (k1, 10 min), k4, the k5 corpus on one GPU, streaming latency (k3), the CPU baseline.

N > 1 (BASELINE config 5, "k5"): a step = the whole corpus build of prepare_corpus.py:179-236 on 64 sessions x 128 ch x 1 h
dealt round-robin to the ranks: per-day baseline statistics from the baseline sessions, ONE NCCL all-gather of all days'
(n, mean, M2) + fixed-order merge, then every owned session extracted and z-scored.  Everything is inside the timed region;
strong scaling (the corpus is fixed).  Parity (mu / sigma identical on every rank and equal to numpy over the baseline
features; one session's features equal to a single-GPU recomputation) is asserted outside the timed region (`parity_ok`).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {  # name: (fs, channels, seconds)
    "k1": (1000, 128, 60),
    "m10": (1000, 128, 600),
    "k2": (1000, 128, 3600),
    "k4": (2000, 1024, 600),
}
K5_SESSIONS, K5_DAYS = 64, 8          # 8 days x 8 sessions; session 0 of each day is its baseline
METRIC = "hga_channel_samples_per_s"
UNIT = "channel-samples/s"
TOLERANCE = {"f64": "max-abs 1e-9 per feature vs the float64 oracle", "f32": "max-abs 1e-4 per feature (5e-4 at 2 kHz)",
             "mixed": "max-abs 5e-5 per feature (2e-4 at 2 kHz)"}


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def workload_name(w):
    fs, C, secs = WORKLOADS[w]
    return f"{w}: {C} ch x {secs} s @ {fs} Hz float64, band-pass+notch cascade, 50ms/10ms log-power, z-score"


def k5_name():
    return (f"k5: {K5_SESSIONS} sessions x 128 ch x 3600 s @ 1000 Hz float64 ({K5_DAYS} days, session 0 of a day = its baseline), per-day "
            "baseline statistics + all-gather + fixed-order merge + extraction + z-score of every session")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 7:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own CPU composition (scipy sosfilt x2 -> compiled reference extension
# compute_log_power_features -> numpy z-score, local/units.py:151-160), see oracle/.  The composition is single-threaded
# and channels never mix, so "all host cores" = one worker process per block of channels.  A worker builds its block ONCE
# (pool initializer) and returns a checksum per step: nothing but an index and a float crosses the process boundary inside
# the timed region (round 1 pickled the 7.7 MB blocks on every step and measured the pipe, not the filters).
# ----------------------------------------------------------------------------------------------------------------------
def _host_cores():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def _ref_worker(k, fs, n_rows, cols, seed, start, done, stop, sums):
    """Worker k owns channel block `cols` for its whole life: generate once, then one extraction per start/done barrier pair."""
    from oracle import build_ref, hga_oracle as O
    ref = build_ref.load()
    log_power = ref.compute_log_power_features if ref is not None else O.compute_log_power_features_fast
    rng = np.random.Generator(np.random.Philox(key=[seed, int(cols[0])]))
    x = rng.standard_normal((n_rows, len(cols)))
    x += (5.0 * np.sin(2 * np.pi * 120.0 * np.arange(n_rows) / fs))[:, None]
    while True:
        start.wait()
        if stop.is_set():
            return
        f = O.HighGammaExtractor(fs, x.shape[1], log_power=log_power).extract_features(x)     # local/units.py:151-156
        sums[k] = float(((f - 3.0) / 0.4).sum())                                                # prepare_corpus.py:226-227
        done.wait()


class ReferencePool:
    """`procs` processes, each holding a block of channels of a (n_rows, C) recording."""

    def __init__(self, fs, C, n_rows, procs):
        import multiprocessing as mp
        from oracle import build_ref
        self.kind = "reference" if build_ref.load() is not None else "port"
        ctx = mp.get_context("fork")
        self.procs = max(1, min(procs, C))
        blocks = np.array_split(np.arange(C), self.procs)
        self.start, self.done, self.stop = ctx.Barrier(self.procs + 1), ctx.Barrier(self.procs + 1), ctx.Event()
        self.sums = ctx.Array("d", self.procs)
        self.ps = [ctx.Process(target=_ref_worker, args=(k, fs, n_rows, blocks[k], 20241104, self.start, self.done, self.stop, self.sums), daemon=True)
                   for k in range(self.procs)]
        for p in self.ps:
            p.start()
        self.units = n_rows * C

    def step(self):
        """One pass of every worker over its block; returns the wall time between the two barriers."""
        self.start.wait()
        t0 = time.perf_counter()
        self.done.wait()
        return time.perf_counter() - t0

    def close(self):
        self.stop.set()
        self.start.wait()
        for p in self.ps:
            p.join(timeout=10)


def reference_rate(fs, C, seconds, procs, steps, warm=1):
    """channel-samples/s of the reference composition: median over `steps` passes of `procs` workers over a (seconds * fs, C)
    recording (channels split across the workers)."""
    pool = ReferencePool(fs, C, fs * seconds, procs)
    try:
        for _ in range(warm):
            pool.step()
        times = [pool.step() for _ in range(steps)]
    finally:
        pool.close()
    return pool.units / float(np.median(times)), pool.kind, times


def run_reference(args):
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    fs, C, secs = WORKLOADS["k2"]
    cores = _host_cores()
    # N = 1: every step is one whole k2 recording (the GPU arm's step).  N > 1: the GPU arm's step is the 64-session corpus
    # (64 x k2 = 29.5 G channel-samples, a minute of CPU time per step on this host): a step here is ONE session of it.
    n_rows = fs * secs
    pool = ReferencePool(fs, C, n_rows, cores)
    procs, kind = pool.procs, pool.kind
    try:
        for _ in range(max(1, min(args.warmup, 2))):
            pool.step()
        times = [pool.step() for _ in range(args.steps)]
    finally:
        pool.close()
    ms = 1e3 * float(np.mean(times))
    value = n_rows * C / (ms / 1e3)
    one_core, _, _ = reference_rate(fs, max(1, C // procs), 60, 1, steps=3)   # one worker's share of the channels, 60 s, one core
    sample = (f"{C} ch x {secs} s @ {fs} Hz white float64 per step (one whole k2 recording" +
              ("" if world == 1 else f"; stands for one of the {K5_SESSIONS} sessions of k5") +
              f"), channels split over {procs} worker processes that hold their block")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak" if world == 1 else "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name("k2") if world == 1 else k5_name(), "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": kind, "sample": sample,
                         "one_core_value": one_core, "parallel_efficiency": value / (procs * one_core)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ----------------------------------------------------------------------------------------------------------------------
# B200 arm, shared pieces
# ----------------------------------------------------------------------------------------------------------------------
def make_recording(torch, dev, T, C, fs, seed):
    """Synthetic recording generated on the device (Philox): white N(0, 1) + a 120 Hz line sitting in the notch."""
    g = torch.Generator(device=dev)
    g.manual_seed(int(seed))
    x = torch.empty((T, C), dtype=torch.float64, device=dev)
    step_rows = max(1, (1 << 27) // C)
    for r0 in range(0, T, step_rows):
        x[r0:r0 + step_rows].normal_(generator=g)
    x += (5.0 * torch.sin(2 * torch.pi * 120.0 * torch.arange(T, device=dev, dtype=torch.float64) / fs)).unsqueeze(1)
    return x


def time_steps(torch, fn, steps, warm, barrier=None):
    """ms per call of fn(): `warm` untimed calls, then `steps` calls between two CUDA events on the current stream."""
    for _ in range(warm):
        fn()
    if barrier:
        barrier()
    else:
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    if barrier:
        barrier()
    else:
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def roofline_entry(C, T, hop, ms, hbm_peak):
    bytes_per = 8.0 + 8.0 / hop
    achieved = C * T * bytes_per / (ms / 1e3) / 1e9
    return achieved, achieved / hbm_peak, bytes_per


def profile_facts(precision):
    """ncu facts of the committed capture of this round (profiles/r02_ncu_facts.json): DRAM traffic per launch and pipe
    utilisation of the dominant kernel.  Not measured here (nothing in this file runs under a profiler)."""
    p = ROOT / "profiles" / "r02_ncu_facts.json"
    if p.exists():
        return json.loads(p.read_text()).get(precision)
    return None


def bind_near_gpu(torch, local):
    """Pin this process to the CPUs next to its GPU (first-touch then places pinned staging on that NUMA node)."""
    try:
        out = subprocess.run(["nvidia-smi", f"--id={local}", "--query-gpu=pci.bus_id", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
        bus = out[-12:].lower() if out else None          # 00000000:1B:00.0 -> 0000:1b:00.0 (sysfs name)
        cpus = Path(f"/sys/bus/pci/devices/{str(bus).lower()}/local_cpulist").read_text().strip()
        node = Path(f"/sys/bus/pci/devices/{str(bus).lower()}/numa_node").read_text().strip()
        ids = set()
        for part in cpus.split(","):
            a, _, b = part.partition("-")
            ids.update(range(int(a), int(b or a) + 1))
        ids &= os.sched_getaffinity(0)
        if ids:
            os.sched_setaffinity(0, ids)
        return {"pci": str(bus), "numa_node": node, "cpus": cpus, "bound": bool(ids)}
    except Exception as exc:      # report, do not hide
        return {"error": repr(exc)}


def h2d_bandwidth(torch, dev, barrier=None, mib=512, reps=3):
    """GB/s of a bare pinned host -> device copy on this rank (all ranks at once when a barrier is given)."""
    h = torch.empty(mib << 20, dtype=torch.uint8, pin_memory=True)
    d = torch.empty(mib << 20, dtype=torch.uint8, device=dev)
    d.copy_(h, non_blocking=True)
    ms = time_steps(torch, lambda: d.copy_(h, non_blocking=True), reps, 1, barrier)
    return (mib << 20) / (ms / 1e3) / 1e9


def stream_latency(fs, C, mu, sigma, device, n_push=10500, warm=500, chunk=10):
    """p50 / p99 per-packet latency (us) of hga_stream_push_host on ONE long-lived handle (the packet's way to the device,
    one kernel and the frame's way back inside the timed call), of the wire-packet call, and of the reference composition fed
    the same packets (bounded: 2000 packets)."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    from delayed_speech_synthesis_b200 import wire
    x = np.random.default_rng(20241104).standard_normal(((n_push + warm) * chunk, C))
    eng = HgaEngine(fs, C, precision="exact", device=device)
    lat = np.empty(n_push + warm)
    for i in range(n_push + warm):
        pkt = x[i * chunk:(i + 1) * chunk]
        t0 = time.perf_counter()
        eng.push(pkt, mu, sigma)
        lat[i] = time.perf_counter() - t0
    lat = lat[warm:] * 1e6
    n_pk = min(n_push, 4000)
    x32 = x[:(n_pk + warm) * chunk].astype(np.float32)
    pkts = [wire.pack_packet(x32[i * chunk:(i + 1) * chunk].T) for i in range(n_pk + warm)]
    eng_p = HgaEngine(fs, C, precision="exact", device=device)
    lp = np.empty(n_pk + warm)
    for i, pk in enumerate(pkts):
        t0 = time.perf_counter()
        eng_p.push_packet(pk, mu, sigma)
        lp[i] = time.perf_counter() - t0
    lp = lp[warm:] * 1e6
    from oracle import build_ref, hga_oracle as O   # cpu-baseline leg: the reference composition on the same packets
    ref = build_ref.load()
    oex = O.HighGammaExtractor(fs, C, post_transforms=[O.ZScoreNormalization(mu, sigma)],
                               log_power=(ref.compute_log_power_features if ref is not None else O.compute_log_power_features_fast))
    n_ref = 2000
    rl = np.empty(n_ref)
    for i in range(n_ref):
        pkt = x[i * chunk:(i + 1) * chunk]
        t0 = time.perf_counter()
        oex.extract_features(pkt)
        rl[i] = time.perf_counter() - t0
    rl = rl[200:] * 1e6
    return {"packet": f"{chunk} x {C} float64", "n": int(n_push), "p50_us": float(np.percentile(lat, 50)),
            "p99_us": float(np.percentile(lat, 99)), "max_us": float(lat.max()),
            "path": "hga_stream_push_host on one long-lived handle: mapped pinned staging (zero-copy), 1 kernel (k_stream_push_sys), completion slots polled in mapped memory, exact arithmetic",
            "wire_packet": {"packet": f"7-byte header + {C} x {chunk} float32", "n": int(n_pk), "p50_us": float(np.percentile(lp, 50)),
                            "p99_us": float(np.percentile(lp, 99)), "path": "hga_stream_push_packet (interpret_bytes fused into the kernel)"},
            "reference_p50_us": float(np.percentile(rl, 50)), "reference_p99_us": float(np.percentile(rl, 99)),
            "reference_n": int(len(rl))}


def device_rate(torch, eng, x, mu, sigma, steps, warm=3):
    """(ms per extraction, features) of the device-resident call on x."""
    out = torch.empty((eng.num_frames(x.shape[0]), x.shape[1]), dtype=torch.float64, device=x.device)
    ms = time_steps(torch, lambda: eng.extract_offline(x, None, mu, sigma, out=out), steps, warm)
    return ms, out


def chain_side(torch, dev, local, steps, hbm_peak):
    """K2 through the reference's pre-transform chain (prepare_corpus.py:152-160: select both grids -> per-grid common average
    -> select the speech grid; 129 raw columns -> 64 channels), packed float32: the chain inside the extraction kernel's load
    stage (hga_extract_offline_mapped) against the same two launches as separate passes."""
    from delayed_speech_synthesis_b200 import transforms as TR
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, c_raw, C, secs = 1000, 129, 64, 3600
    T = fs * secs
    x = make_recording(torch, dev, T, c_raw, fs, 4242)
    x += torch.linspace(-500.0, 500.0, c_raw, dtype=torch.float64, device=dev)     # per-channel offsets, as raw counts x gain have
    layout = np.arange(128) + 1
    grids = [np.flip(layout[:64].reshape(8, 8), axis=0), layout[64:].reshape(8, 8)]
    pre = [TR.ChannelSelection(np.arange(128)), TR.CommonAverageReferencing([19, 38, 77], grids, layout), TR.ChannelSelection(np.arange(64, 128))]
    cmap = TR.ChannelMap(TR.fold(pre, c_raw), local)
    eng = HgaEngine(fs, C, precision="f32", device=local)
    F = eng.num_frames(T)
    out = torch.empty((F, C), dtype=torch.float64, device=dev)
    ms_f = time_steps(torch, lambda: eng.extract_offline_mapped(x, cmap, None, out=out), steps, 3)
    fused = bool(eng.last_fused)
    eng.set_timing(True)
    eng.extract_offline_mapped(x, cmap, None, out=out)
    main_ms = eng.last_kernel_split_ms()[0] if fused else None
    eng.set_timing(False)
    xm = torch.empty((T, C), dtype=torch.float64, device=dev)
    out2 = torch.empty_like(out)

    def two():
        cmap.apply(x, out=xm)
        eng.extract_offline(xm, None, out=out2)
    ms_2 = time_steps(torch, two, steps, 3)
    ms_map = time_steps(torch, lambda: cmap.apply(x, out=xm), steps, 1)
    raw_bytes, out_bytes = T * c_raw * 8, F * C * 8
    res = {"workload": f"k2 through the pre-transform chain: {c_raw} raw columns -> {C} channels x {secs} s @ {fs} Hz, float32 arithmetic",
           "fused": fused, "ms_per_step": ms_f, "value": C * T / (ms_f / 1e3), "main_kernel_ms": main_ms,
           "bytes_per_step": raw_bytes + out_bytes, "achieved_gbs": (raw_bytes + out_bytes) / (ms_f / 1e3) / 1e9,
           "frac_hbm": (raw_bytes + out_bytes) / (ms_f / 1e3) / 1e9 / hbm_peak,
           "two_pass": {"ms_per_step": ms_2, "chanmap_ms": ms_map, "value": C * T / (ms_2 / 1e3),
                        "bytes_per_step": raw_bytes + 2 * T * C * 8 + out_bytes},
           "max_abs_diff_fused_vs_two_pass": float((out - out2).abs().max().item()), "plan": eng.last_plan()}
    del x, xm, out, out2, eng, cmap
    torch.cuda.empty_cache()
    return res


def side_workload(torch, dev, local, name, precisions, steps, hbm_peak):
    """A further BASELINE configuration on the same GPU (short recordings, k4): channel-samples/s per precision."""
    from delayed_speech_synthesis_b200.engine import HgaEngine
    fs, C, secs = WORKLOADS[name]
    T = fs * secs
    x = make_recording(torch, dev, T, C, fs, 777)
    res = {"workload": workload_name(name), "input_bytes": T * C * 8,
           "l2": "input > 126 MB L2" if T * C * 8 > (126 << 20) else "input fits the 126 MB L2 (recurring steps re-read it from L2)"}
    for prec in precisions:
        eng = HgaEngine(fs, C, precision=prec, device=local)
        base = eng.extract_offline(x[: 60 * fs])
        mu, sigma = base.mean(0), base.std(0, unbiased=False)
        ms, _ = device_rate(torch, eng, x, mu, sigma, steps)
        ach, frac, _ = roofline_entry(C, T, eng.hop, ms, hbm_peak)
        res[prec] = {"value": C * T / (ms / 1e3), "ms_per_step": ms, "roofline_frac": frac, "plan": eng.last_plan()}
        del eng
    del x
    torch.cuda.empty_cache()
    return res


# ----------------------------------------------------------------------------------------------------------------------
# k5: the corpus build (prepare_corpus.py:179-236) on 64 sessions, sharded over the ranks
# ----------------------------------------------------------------------------------------------------------------------
class Corpus:
    """The rank's share of k5.  Sessions are generated on the device from their id (so any rank can regenerate any session
    for the parity check) and stay resident; a rank that owns more sessions than fit beside the scratch (N = 1: 64 x 3.69 GB
    = 236 GB > 180 GB of HBM) keeps `n_buf` distinct recordings and session i of its list reads recording i % n_buf: every
    extraction still streams 3.69 GB from HBM, only the content repeats."""

    def __init__(self, torch, dev, local, rank, world, precision, max_buffers=32):
        from delayed_speech_synthesis_b200 import corpus
        from delayed_speech_synthesis_b200.engine import HgaEngine
        from delayed_speech_synthesis_b200.ops import ChannelStats
        self.torch, self.dev, self.local, self.rank, self.world = torch, dev, local, rank, world
        self.fs, self.C, secs = WORKLOADS["k2"]
        self.T = self.fs * secs
        self.corpus, self.ChannelStats = corpus, ChannelStats
        self.day_of = [s // (K5_SESSIONS // K5_DAYS) for s in range(K5_SESSIONS)]
        self.baseline = [d * (K5_SESSIONS // K5_DAYS) for d in range(K5_DAYS)]
        self.mine = corpus.shard_weighted([2 if s in self.baseline else 1 for s in range(K5_SESSIONS)], rank, world)
        free, _ = torch.cuda.mem_get_info(dev)
        per = self.T * self.C * 8
        self.n_buf = max(1, min(len(self.mine), max_buffers, int((free - (14 << 30)) // per)))
        self.buffers = [make_recording(torch, dev, self.T, self.C, self.fs, 5000 + s) for s in self.mine[:self.n_buf]]
        self.eng = HgaEngine(self.fs, self.C, precision=precision, device=local)
        self.out = torch.empty((self.eng.num_frames(self.T), self.C), dtype=torch.float64, device=dev)
        self.units = K5_SESSIONS * self.C * self.T        # channel-samples of the corpus (the baseline pass is overhead)

    def session(self, s):
        return self.buffers[self.mine.index(s) % self.n_buf]

    def extract(self, x, mu, sigma):
        return self.eng.extract_offline(x, None, mu, sigma, out=self.out)

    def accumulate(self, f):
        return self.ChannelStats(self.C, self.local).update(f).acc

    def extract_stats(self, x):
        st = self.ChannelStats(self.C, self.local)
        self.eng.extract_offline_stats(x, st, None, out=self.out)
        return st.acc

    def step(self, timings=None, on_features=None):
        return self.corpus.extract_corpus(K5_SESSIONS, self.day_of, self.baseline, self.session, self.extract, self.accumulate,
                                          n_channels=self.C, device=self.local, on_features=on_features, timings=timings,
                                          extract_stats=self.extract_stats)

    # ---- parity, outside the timed region
    def parity(self, dist):
        torch = self.torch
        mu, sigma, slots = self.step()
        ok, detail = True, {}
        if self.world > 1:       # (1) bit-identical mu / sigma on every rank
            both = torch.stack([mu, sigma]).contiguous()
            allr = [torch.empty_like(both) for _ in range(self.world)]
            dist.all_gather(allr, both)
            detail["mu_sigma_identical_on_all_ranks"] = all(bool(torch.equal(allr[0], a)) for a in allr)
            ok &= detail["mu_sigma_identical_on_all_ranks"]
        worst = 0.0              # (2) against numpy over the baseline features of the days this rank owns the baseline of
        for s in self.baseline:
            if s in self.mine and self.mine.index(s) < self.n_buf:
                f = self.extract(self.session(s), None, None).cpu().numpy()
                d = slots[self.day_of[s]]
                worst = max(worst, float(np.abs(mu[d].cpu().numpy() / f.mean(0) - 1).max()), float(np.abs(sigma[d].cpu().numpy() / f.std(0) - 1).max()))
        t = torch.tensor([worst], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        detail["mu_sigma_max_rel_err_vs_numpy"] = float(t.item())
        ok &= detail["mu_sigma_max_rel_err_vs_numpy"] <= 1e-12
        # (3) one session of another rank, recomputed on rank 0 from its id with the merged statistics
        owners = {s: r for r in range(self.world) for s in self.corpus.shard_weighted([2 if s in self.baseline else 1 for s in range(K5_SESSIONS)], r, self.world)}
        s_chk = next((s for s in range(K5_SESSIONS) if owners[s] != 0), 1)       # a session rank 0 does not own (any, at N = 1)
        owner = owners[s_chk]
        d = slots[self.day_of[s_chk]]
        feats = torch.empty_like(self.out)
        if self.rank == owner:
            feats.copy_(self.extract(self.session(s_chk), mu[d], sigma[d]))
        if self.world > 1:
            dist.broadcast(feats, src=owner)
        same = True
        if self.rank == 0:
            x = self.session(s_chk) if owner == 0 else make_recording(torch, self.dev, self.T, self.C, self.fs, 5000 + s_chk)
            mine = self.extract(x, mu[d], sigma[d])
            same = bool(torch.equal(mine, feats)) and bool(torch.isfinite(mine).all())
            del x
        t = torch.tensor([1.0 if same else 0.0], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.broadcast(t, src=0)
        detail["session_features_equal_single_gpu"] = bool(t.item() == 1.0)
        ok &= detail["session_features_equal_single_gpu"]
        detail["checked_session"] = s_chk
        return bool(ok), detail


def corpus_timed(torch, cp, steps, warm, barrier):
    """(ms per corpus step on this rank, mean ms of the statistics exchange inside it)."""
    timings = {}
    for _ in range(warm):
        cp.step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        cp.step(timings)
    e1.record()
    barrier()
    ev = timings.get("exchange_events", [])
    exch = float(np.mean([a.elapsed_time(b) for a, b in ev])) if ev else None
    return e0.elapsed_time(e1) / steps, exch


def corpus_e2e(torch, cp, dist, steps, barrier):
    """The same corpus step through the host-buffer C-ABI call: every session goes up from pinned host memory and its
    features come down, inside the timed region.  One host recording per rank stands for all of its sessions (same bytes
    through PCIe every time).  Returns (seconds per step, max over ranks; bytes up / down per step on this rank)."""
    xh = torch.empty((cp.T, cp.C), dtype=torch.float64, pin_memory=True)
    xh.copy_(cp.buffers[0])
    oh = torch.empty((cp.eng.num_frames(cp.T), cp.C), dtype=torch.float64, pin_memory=True)
    xn, on = xh.numpy(), oh.numpy()

    def extract(x, mu, sigma):
        m = None if mu is None else mu.cpu().numpy()
        s = None if sigma is None else sigma.cpu().numpy()
        return cp.eng.extract_offline(xn, None, m, s, out=on)

    def step():
        cp.corpus.extract_corpus(K5_SESSIONS, cp.day_of, cp.baseline, lambda s: xn, extract, cp.accumulate, n_channels=cp.C, device=cp.local)

    cp.eng.extract_offline(xn, None, None, None, out=on)     # staging buffers, plan
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    t = torch.tensor([dt], dtype=torch.float64, device=cp.dev)
    if cp.world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    n_ext = len(cp.mine) + sum(1 for s in cp.baseline if s in cp.mine)
    del xh, oh
    return float(t.item()), n_ext * cp.T * cp.C * 8, n_ext * cp.eng.num_frames(cp.T) * cp.C * 8


def raw_int16_e2e(torch, eng, x, mu_h, sg_h, n_calls, dist, world, barrier):
    """hga_extract_offline_raw_host on int16 amplifier counts + per-channel gain (BCI2000MatFile.signals() formed on the
    device): a quarter of the upload bytes.  Side entry: the input differs from the headline's."""
    from delayed_speech_synthesis_b200 import transforms as TR
    T, C = x.shape
    gain_h = np.linspace(0.05, 0.15, C)
    counts = torch.clamp(torch.round(x / torch.from_numpy(gain_h).to(x.device)), -32000, 32000).to(torch.int16)
    ch = torch.empty((T, C), dtype=torch.int16, pin_memory=True)
    ch.copy_(counts)
    oh = torch.empty((eng.num_frames(T), C), dtype=torch.float64, pin_memory=True)
    ident = TR.ChannelMap(TR.ChannelMapSpec.identity(C), eng.device)
    cn, on = ch.numpy(), oh.numpy()
    eng.extract_offline_raw(cn, ident, gain_h, None, mu_h, sg_h, out=on)
    barrier()
    t0 = time.perf_counter()
    for _ in range(n_calls):
        eng.extract_offline_raw(cn, ident, gain_h, None, mu_h, sg_h, out=on)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n_calls
    t = torch.tensor([dt], dtype=torch.float64, device=x.device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    want = eng.extract_offline(ident.apply_raw(counts, gain_h), None, torch.from_numpy(mu_h).to(x.device), torch.from_numpy(sg_h).to(x.device))
    ok = bool(torch.equal(oh.to(x.device), want))
    del ch, oh, counts, want
    return {"value": world * C * T / float(t.item()), "unit": UNIT, "ms_per_recording": 1e3 * float(t.item()), "h2d_bytes_per_recording": int(T * C * 2),
            "matches_device_result": ok,
            "path": "hga_extract_offline_raw_host: int16 counts x float64 gain expanded on the device (k_rawscale<int16, gain>), one recording per rank"}


# ----------------------------------------------------------------------------------------------------------------------
def run_single(args):
    import torch

    from delayed_speech_synthesis_b200 import _lib
    from delayed_speech_synthesis_b200.engine import HgaEngine

    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the HGA path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    lib = _lib.load()
    fs, C, secs = WORKLOADS[args.workload]
    T = fs * secs
    hbm_peak, peak_src = peaks()
    placement = bind_near_gpu(torch, local)

    x = make_recording(torch, dev, T, C, fs, 20241104)
    eng = HgaEngine(fs, C, precision=args.precision, device=local)
    base = eng.extract_offline(x[: 60 * fs])                    # day baseline from the first 60 s
    mu, sigma = base.mean(0), base.std(0, unbiased=False)
    F = eng.num_frames(T)
    out = torch.empty((F, C), dtype=torch.float64, device=dev)

    def step():
        eng.extract_offline(x, None, mu, sigma, out=out)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    # launch-sequence time (main kernel + carry fix-up): CUDA events recorded by the library on the launching stream
    eng.set_timing(True)
    kms, main_ms = [], []
    for _ in range(min(args.steps, 10)):
        step()
        kms.append(eng.last_kernel_ms())
        main_ms.append(eng.last_kernel_split_ms()[0])
    eng.set_timing(False)
    kernel_ms = float(np.median(kms))
    main_kernel_ms = float(np.median(main_ms))
    l0 = lib.hga_launch_count()
    with ClockSampler(local) as clk:
        ms = time_steps(torch, step, args.steps, 0)
    launches = lib.hga_launch_count() - l0
    value = C * T / (ms / 1e3)
    plan = eng.last_plan()
    achieved, frac, bytes_per = roofline_entry(C, T, eng.hop, kernel_ms, hbm_peak)

    # ---- the other precisions on the same resident input (same timing rules)
    other = {}
    for prec in ("f64", "f32", "mixed"):
        if prec == args.precision:
            continue
        try:
            e2 = HgaEngine(fs, C, precision=prec, device=local)
            ms2, o2 = device_rate(torch, e2, x, mu, sigma, args.steps)
            other[prec] = {"value": C * T / (ms2 / 1e3), "ms_per_step": ms2, "roofline_frac": roofline_entry(C, T, e2.hop, ms2, hbm_peak)[1],
                           "max_abs_diff_vs_headline": float((o2 - out).abs().max().item()), "tolerance": TOLERANCE[prec], "plan": e2.last_plan(),
                           "ncu": profile_facts(prec)}
            del e2, o2
        except Exception as exc:
            other[prec] = {"error": repr(exc)}

    # ---- end to end through the host-buffer C-ABI call: host input, host output, copies inside the timed region
    e2e = None
    try:
        if args.no_e2e:
            raise RuntimeError("skipped (--no-e2e)")
        h2d = h2d_bandwidth(torch, dev)
        xh = torch.empty((T, C), dtype=torch.float64, pin_memory=True)
        xh.copy_(x)
        oh = torch.empty((F, C), dtype=torch.float64, pin_memory=True)
        mu_h, sg_h = mu.cpu().numpy(), sigma.cpu().numpy()
        xn, on = xh.numpy(), oh.numpy()
        n_e2e = max(2, min(args.steps, 5))
        eng.extract_offline(xn, None, mu_h, sg_h, out=on)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            eng.extract_offline(xn, None, mu_h, sg_h, out=on)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / n_e2e
        e2e = {"value": C * T / dt, "unit": UNIT, "h2d_bytes_per_step": int(T * C * 8), "d2h_bytes_per_step": int(F * C * 8),
               "ms_per_step": 1e3 * dt, "matches_device_result": bool(torch.equal(oh.to(dev), out)), "host_buffers": "page-locked (read by the DMA engine where they lie)",
               "h2d_gbs_bare_copy": h2d, "placement": placement}
        # pageable buffers (what BCI2000MatFile.signals() hands over): staged through the library's pinned ping-pong buffers
        xp = np.empty((T, C), dtype=np.float64)
        xp[:] = xn
        op = np.empty((F, C), dtype=np.float64)
        eng.extract_offline(xp, None, mu_h, sg_h, out=op)
        t0 = time.perf_counter()
        for _ in range(2):
            eng.extract_offline(xp, None, mu_h, sg_h, out=op)
        dtp = (time.perf_counter() - t0) / 2
        e2e["pageable"] = {"value": C * T / dtp, "ms_per_step": 1e3 * dtp, "matches_device_result": bool(np.array_equal(op, on)),
                           "path": "pageable numpy in / out: host copies into two 64 MB pinned staging buffers overlap the DMA"}
        del xp, op, xh
        try:
            e2e["raw_int16"] = raw_int16_e2e(torch, eng, x, mu_h, sg_h, n_e2e, None, 1, torch.cuda.synchronize)
        except Exception as exc:
            e2e["raw_int16"] = {"error": repr(exc)}
        del oh
    except Exception as exc:  # report, do not hide
        e2e = {"value": None, "unit": UNIT, "error": repr(exc)}

    # ---- streaming (BASELINE config 3): 10-sample x C packets on one long-lived handle
    stream = None
    if not args.no_stream:
        try:
            stream = stream_latency(fs, C, mu.cpu().numpy(), sigma.cpu().numpy(), local)
        except Exception as exc:
            stream = {"error": repr(exc)}
    del x, out
    torch.cuda.empty_cache()

    # ---- other BASELINE configurations on this GPU: short recordings (the chunking's worst case), k4, the k5 corpus
    sides = {}
    if not args.no_sides:
        for name, precs, st in (("k1", ("f32", "f64"), 20), ("m10", ("f32", "f64"), 20), ("k4", ("f32", "f64"), 10)):
            try:
                sides[name] = side_workload(torch, dev, local, name, precs, st, hbm_peak)
            except Exception as exc:
                sides[name] = {"error": repr(exc)}
        try:
            sides["k2_chain"] = chain_side(torch, dev, local, 10, hbm_peak)
        except Exception as exc:
            sides["k2_chain"] = {"error": repr(exc)}
        try:
            cp = Corpus(torch, dev, local, 0, 1, args.precision)
            ms5, exch = corpus_timed(torch, cp, 3, 1, torch.cuda.synchronize)
            ok5, det5 = cp.parity(None)
            sides["k5"] = {"workload": k5_name(), "value": cp.units / (ms5 / 1e3), "ms_per_step": ms5, "stats_exchange_ms": exch, "parity_ok": ok5, "parity": det5,
                           "resident_recordings": cp.n_buf, "note": "64 sessions on one GPU read 32 resident recordings twice (236 GB of input, 180 GB of HBM)"}
            del cp
        except Exception as exc:
            sides["k5"] = {"error": repr(exc)}
        torch.cuda.empty_cache()

    cpu = None
    if not args.no_cpu_baseline:
        cores = _host_cores()
        v1, kind, _ = reference_rate(fs, C, 60, 1, steps=3)                    # the composition as the reference runs it: one process
        vn, _, _ = reference_rate(fs, C, 600, cores, steps=2)                  # channels split over all host cores, 10 min sample
        cpu = {"value": vn, "unit": UNIT, "cores": min(cores, C), "kind": kind,
               "sample": f"{C} ch x 600 s @ {fs} Hz white float64, channels split over {min(cores, C)} processes that hold their block, median of 2 passes",
               "one_core": {"value": v1, "sample": f"{C} ch x 60 s, one process, median of 3"}}
    facts = profile_facts(args.precision)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": args.precision,
        "data": "synthetic",
        "config": {"workload": workload_name(args.workload), "precision": args.precision, "tolerance": TOLERANCE[args.precision],
                   "l2": f"input {T * C * 8 / 1e9:.2f} GB per step > 126 MB L2 (no flush needed)", "plan": plan},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": frac,
                     "traffic": (facts or {}).get("traffic_bytes_per_step"), "peak_source": peak_src, "bytes_per_channel_sample": bytes_per,
                     "kernel": "k_hga_lanes (+ k_lane_fix, k_lane_fix_frames: the carried heads); kernel_ms is the whole launch sequence of a step",
                     "kernel_ms": kernel_ms,
                     "dominant_kernel_alone": {"kernel": "k_hga_lanes", "ms": main_kernel_ms, "achieved": roofline_entry(C, T, eng.hop, main_kernel_ms, hbm_peak)[0],
                                               "frac": roofline_entry(C, T, eng.hop, main_kernel_ms, hbm_peak)[1],
                                               "note": "the main kernel between its own CUDA events; `achieved`/`frac` above charge the same algorithmic bytes to the whole "
                                                       "launch sequence (carry + head fix-up kernels included), which is the lower figure"},
                     "binding_pipe": "fp64" if args.precision == "f64" else "fma (packed f32x2)",
                     "pipe_frac": (facts or {}).get("pipe_frac"), "ncu": facts},
        "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk.summary(),
        "other_precisions": other, "tolerance": TOLERANCE, "sides": sides, "stream_latency": stream,
    }
    print(json.dumps(line))


def run_corpus(args):
    import torch
    import torch.distributed as dist

    from delayed_speech_synthesis_b200 import _lib

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the HGA path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    hbm_peak, peak_src = peaks()
    placement = bind_near_gpu(torch, local)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    cp = Corpus(torch, dev, local, rank, world, args.precision)
    cp.step()                                   # plans, scratch, the communicator's first collective
    parity_ok, parity = cp.parity(dist)
    barrier()
    l0 = lib.hga_launch_count()
    with ClockSampler(local) as clk:
        ms, exch = corpus_timed(torch, cp, args.steps, max(args.warmup, 3), barrier)
    launches = lib.hga_launch_count() - l0
    t = torch.tensor([ms, exch or 0.0], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tmin = torch.tensor([exch or 0.0], dtype=torch.float64, device=dev)
    dist.all_reduce(tmin, op=dist.ReduceOp.MIN)       # the last rank to arrive sees the exchange itself, the others also wait for it
    ms, exch_max, exch = float(t[0].item()), float(t[1].item()), float(tmin.item())
    value = cp.units / (ms / 1e3)
    # per-rank kernel view: one session's launch sequence
    cp.eng.set_timing(True)
    kms = []
    for _ in range(5):
        cp.extract(cp.buffers[0], None, None)
        kms.append(cp.eng.last_kernel_ms())
    cp.eng.set_timing(False)
    kernel_ms = float(np.median(kms))
    achieved, frac, bytes_per = roofline_entry(cp.C, cp.T, cp.eng.hop, kernel_ms, hbm_peak)

    e2e = None
    try:
        if args.no_e2e:
            raise RuntimeError("skipped (--no-e2e)")
        h2d = h2d_bandwidth(torch, dev, barrier)
        hb = torch.tensor([h2d], dtype=torch.float64, device=dev)
        allh = [torch.empty_like(hb) for _ in range(world)]
        dist.all_gather(allh, hb)
        dt, up, down = corpus_e2e(torch, cp, dist, 2, barrier)
        e2e = {"value": cp.units / dt, "unit": UNIT, "h2d_bytes_per_step": int(up), "d2h_bytes_per_step": int(down), "ms_per_step": 1e3 * dt,
               "host_buffers": "page-locked, one recording per rank standing for each of its sessions",
               "h2d_gbs_bare_copy_per_rank_all_ranks_at_once": [float(a.item()) for a in allh], "placement": placement,
               "limiter": "host -> device copies: compare ms_per_step with h2d_bytes_per_step / the bare-copy rate measured with all ranks copying at once"}
        try:
            mu0 = np.full(cp.C, 3.0)
            e2e["raw_int16"] = raw_int16_e2e(torch, cp.eng, cp.buffers[0], mu0, np.full(cp.C, 0.4), 3, dist, world, barrier)
        except Exception as exc:
            e2e["raw_int16"] = {"error": repr(exc)}
    except Exception as exc:
        e2e = {"value": None, "unit": UNIT, "error": repr(exc)}

    if rank == 0:
        facts = profile_facts(args.precision)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": {"workload": k5_name(), "precision": args.precision, "tolerance": TOLERANCE[args.precision], "sessions_per_rank": len(cp.mine),
                       "resident_recordings_per_rank": cp.n_buf, "l2": "every session is 3.69 GB > 126 MB L2 (no flush needed)",
                       "timed_region": "per-day baseline extraction + statistics, NCCL all-gather of all days' (n, mean, M2), fixed-order merge, extraction + z-score of every owned session",
                       "stats_exchange_ms": exch, "stats_exchange_ms_incl_wait_for_slowest_rank": exch_max, "plan": cp.eng.last_plan()},
            "parity_ok": parity_ok, "parity": parity,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": frac, "traffic": (facts or {}).get("traffic_bytes_per_step"),
                         "peak_source": peak_src, "bytes_per_channel_sample": bytes_per, "kernel_ms": kernel_ms,
                         "kernel": "per rank, one session: k_hga_lanes + carried-head fix-up", "pipe_frac": (facts or {}).get("pipe_frac")},
            "cpu_baseline": None, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk.summary(), "tolerance": TOLERANCE,
        }
        print(json.dumps(line))
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="k2", choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default="f32", choices=["f64", "f32", "mixed"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-stream", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sides", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif int(os.environ.get("WORLD_SIZE", "1")) > 1:
        run_corpus(args)
    else:
        run_single(args)


if __name__ == "__main__":
    main()
