"""Aggregate ncu --page source --csv stall samples by SASS address region.  usage: ncu_regions.py src.csv lo:hi:name ..."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
base = int(data[0][0], 16)
regs = [(int(a, 16), int(b, 16), n) for a, b, n in (s.split(':') for s in sys.argv[2:])]
agg = collections.defaultdict(collections.Counter)
for r in data:
    off = int(r[0], 16) - base
    for lo, hi, n in regs:
        if lo <= off < hi:
            agg[n]['samples'] += int(r[ix['# Samples']]); agg[n]['inst'] += int(r[ix['Instructions Executed']])
            for s in stalls: agg[n][s] += int(r[ix[s]])
for lo, hi, n in regs:
    a = agg[n]
    print(f"{n:14s} samples {a['samples']:7d} inst {a['inst']:10d}", {k[6:]: v for k, v in a.items() if k.startswith('stall') and v > 0.03 * max(1, a['samples'])})
