#!/usr/bin/env python3
"""Turn the round's ncu reports (gpurun_out/*.ncu-rep) into the tracked summaries under profiles/.

    python tools/ncu_facts.py r02 f32=gpurun_out/r02_lanes_f32.ncu-rep f64=gpurun_out/r02_lanes_f64.ncu-rep \
           fix_f32=gpurun_out/r02_fix_f32.ncu-rep

writes profiles/<round>_ncu_<key>.csv (the metrics below, one row per captured launch) and merges
profiles/<round>_ncu_facts.json, which bench.py quotes in `roofline.traffic` / `roofline.pipe_frac` (it never runs a profiler).
"""
import csv
import io
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
KEEP = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
    "smsp__average_warp_latency_issue_stalled_wait_per_warp_active.pct", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
]


def to_bytes(value, unit):
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(value.replace(",", "")) * scale[unit]


def to_ms(value, unit):
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "usecond": 1e-3, "msecond": 1.0, "nsecond": 1e-6, "second": 1e3}
    return float(value.replace(",", "")) * scale[unit]


def read_report(path):
    txt = subprocess.run(["ncu", "-i", str(path), "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    head, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        out.append({k: (v, u) for k, v, u in zip(head, r, units)})
    return out


def main():
    rnd = sys.argv[1]
    facts_path = ROOT / "profiles" / f"{rnd}_ncu_facts.json"
    facts = json.loads(facts_path.read_text()) if facts_path.exists() else {}
    for spec in sys.argv[2:]:
        key, _, rep = spec.partition("=")
        launches = read_report(rep)
        dst = ROOT / "profiles" / f"{rnd}_ncu_{key}.csv"
        with dst.open("w", newline="") as fh:
            w = csv.writer(fh)
            w.writerow(["launch", "kernel", "metric", "unit", "value"])
            for i, d in enumerate(launches):
                name = d["Kernel Name"][0].split("(")[0]
                for m in KEEP:
                    if m in d:
                        w.writerow([i, name, m, d[m][1], d[m][0]])
        per = []
        for d in launches:
            rd = to_bytes(*d["dram__bytes_read.sum"])
            wr = to_bytes(*d["dram__bytes_write.sum"])
            fma = float(d["sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed"][0])
            f64 = float(d["sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed"][0])
            per.append({
                "kernel": d["Kernel Name"][0].split("(")[0].replace("void ", ""),
                "grid": d["launch__grid_size"][0], "block": d["launch__block_size"][0], "regs": d["launch__registers_per_thread"][0],
                "ms_under_ncu": to_ms(*d["gpu__time_duration.sum"]),
                "dram_read_bytes": rd, "dram_write_bytes": wr,
                "pipe_fma_pct_elapsed": fma, "pipe_fp64_pct_elapsed": f64,
                "issue_active_pct": float(d["sm__issue_active.avg.pct_of_peak_sustained_elapsed"][0]),
            })
        facts.setdefault("_captures", {})[key] = {"report": Path(rep).name, "launches": per}
    # the per-precision entries bench.py quotes: the dominant kernel's DRAM bytes per launch and its binding pipe
    for prec in ("f32", "f64", "mixed"):
        cap = facts["_captures"].get(prec)
        if not cap:
            continue
        main = cap["launches"][0]
        pipe = "fp64" if prec == "f64" else "fma"
        entry = {
            "kernel": main["kernel"], "report": cap["report"],
            "traffic_bytes_per_step": main["dram_read_bytes"] + main["dram_write_bytes"],
            "dram_read_bytes": main["dram_read_bytes"], "dram_write_bytes": main["dram_write_bytes"],
            "pipe": pipe, "pipe_frac": main[f"pipe_{pipe}_pct_elapsed"] / 100.0, "issue_active_frac": main["issue_active_pct"] / 100.0,
            "ms_under_ncu": main["ms_under_ncu"], "grid": main["grid"], "block": main["block"], "regs": main["regs"],
        }
        fix = facts["_captures"].get(f"fix_{prec}")
        if fix:
            seen, rows = set(), []
            for l in fix["launches"]:   # the first captured launch of each kernel
                if l["kernel"] not in seen:
                    seen.add(l["kernel"])
                    rows.append({k: l[k] for k in ("kernel", "ms_under_ncu", "dram_read_bytes", "dram_write_bytes", "grid", "regs", "pipe_fma_pct_elapsed", "issue_active_pct")})
            entry["fixup_kernels"] = rows
        facts[prec] = entry
    facts_path.write_text(json.dumps(facts, indent=1) + "\n")
    print(facts_path)


if __name__ == "__main__":
    main()
