"""Static issue-cycle estimate of SASS loops: sums the stall field (control bits [105:109) of each 128-bit instruction) over
address ranges of a cuobjdump -sass listing.  usage: cuobjdump -sass bin | python tools/sass_stalls.py <function substring> [lo hi]"""
import re
import sys

want = sys.argv[1]
lo = int(sys.argv[2], 16) if len(sys.argv) > 3 else 0
hi = int(sys.argv[3], 16) if len(sys.argv) > 3 else 1 << 30
lines = sys.stdin.read().splitlines()
infn = False
ins = []
i = 0
while i < len(lines):
    l = lines[i]
    if "Function :" in l:
        infn = want in l
    elif infn:
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);\s+/\* 0x([0-9a-f]{16}) \*/", l)
        if m and i + 1 < len(lines):
            m2 = re.search(r"/\* 0x([0-9a-f]{16}) \*/", lines[i + 1])
            if m2:
                addr = int(m.group(1), 16)
                hiw = int(m2.group(1), 16)
                stall = (hiw >> 41) & 0xF
                yld = (hiw >> 45) & 1
                wbar = (hiw >> 46) & 7
                rbar = (hiw >> 49) & 7
                wait = (hiw >> 52) & 0x3F
                ins.append((addr, m.group(2).strip(), stall, yld, wbar, rbar, wait))
                i += 1
    i += 1
sel = [x for x in ins if lo <= x[0] < hi]
tot = sum(x[2] for x in sel)
from collections import Counter
ops = Counter(x[1].split()[0] if not x[1].startswith("@") else x[1].split()[1] for x in sel)
print(f"{len(sel)} instructions, sum of stall fields = {tot} cycles")
print(ops.most_common(25))
if "-v" in sys.argv:
    for x in sel:
        print(f"{x[0]:05x} st={x[2]:2d} y={x[3]} wb={x[4]} rb={x[5]} wait={x[6]:02x}  {x[1]}")
