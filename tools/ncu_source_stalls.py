#!/usr/bin/env python
"""Top source lines by warp-stall samples from an `ncu --page source --csv --print-source cuda,sass` dump."""
import csv, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
agg = defaultdict(lambda: [0, 0, ""]); cur = ""; hdr = None; seen = []
def num(x):
    try: return int(x)
    except ValueError: return 0
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split('/')[-1]
        if seen and seen[0] == cur: break
        seen.append(cur); continue
    if r and r[0] == "Line No":
        hdr = r; iex = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples"); continue
    if hdr is None or not r[0].isdigit(): continue
    k = (cur, int(r[0])); agg[k][0] += num(r[iex]); agg[k][1] += num(r[isamp]); agg[k][2] = r[1].strip()[:90]
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{k[0]:14s}:{k[1]:4d} samp {100*v[1]/tot:5.1f}% inst {v[0]:8d}  {v[2]}")
