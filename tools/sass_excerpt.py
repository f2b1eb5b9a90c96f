#!/usr/bin/env python3
"""SASS evidence for the lane kernels: opcode histogram plus two excerpts per kernel (the TMA refill + mbarrier hand-off
of the x ring, and the densest stretch of packed-float32 / float64 arithmetic of the filter loop).

    python tools/sass_excerpt.py > profiles/r02_sass_k_hga_lanes.txt
"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "delayed_speech_synthesis_b200" / "libhga_b200.so"
WANT = [("k_hga_lanes<LaneF32, hop 10, TB 60, NI 2, NS 2>", "k_hga_lanesINS_7LaneF32ELi10ELi60ELi2ELi2ELi2ELi3ELi5ELb0ELb0ELb0E", "FFMA2"),
        ("k_hga_lanes<LaneF64, hop 10, TB 80, NI 2, NS 2>", "k_hga_lanesINS_7LaneF64ELi10ELi80E", "DFMA"),
        ("k_lane_fix<float, hop 10>", "k_lane_fixIfLi10E", "FFMA2")]


def functions(text):
    out, name, body = {}, None, []
    for line in text.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            if name:
                out[name] = body
            name, body = m.group(1), []
        elif name:
            m = re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?);", line)
            if m:
                body.append((m.group(1), m.group(2).strip()))
    if name:
        out[name] = body
    return out


def opcode(ins):
    t = ins.split()
    return (t[1] if t[0].startswith("@") else t[0]).split(".")[0]


def main():
    text = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    fns = functions(text)
    print(f"cuobjdump -sass {LIB.relative_to(ROOT)}  (sm_100a)\n")
    for title, key, dense in WANT:
        name = next(n for n in fns if key in n)
        body = fns[name]
        hist = collections.Counter(opcode(i) for _, i in body)
        print("=" * 110)
        print(f"{title}\n{name}\n{len(body)} instructions; " + ", ".join(f"{k} {v}" for k, v in hist.most_common(14)))
        tma = next((i for i, (_, ins) in enumerate(body) if "UTMALDG" in ins and i > 300), None)
        if tma is None:
            tma = next((i for i, (_, ins) in enumerate(body) if "LDGSTS" in ins), None)
        if tma is not None:
            print("\n-- input refill (TMA box / cp.async) and its mbarrier traffic")
            for a, ins in body[max(0, tma - 12): tma + 6]:
                print(f"  /*{a}*/  {ins} ;")
        best, at = -1, 0
        for i in range(0, len(body) - 30):
            n = sum(1 for _, ins in body[i:i + 30] if opcode(ins) in (dense, "FADD2", "DADD", "DFMA", "FFMA2"))
            if n > best:
                best, at = n, i
        print(f"\n-- filter loop, densest 30 instructions ({best} of 30 on the FP pipe)")
        for a, ins in body[at:at + 30]:
            print(f"  /*{a}*/  {ins} ;")
        print()


if __name__ == "__main__":
    sys.exit(main())
