#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by CUDA source line:
instructions executed and stall samples per line.  usage: ncu_source_hot.py dump.csv [top]"""
import csv, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = defaultdict(lambda: [0, 0, ""])
cur_file = ""
seen = []
hdr = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        if cur_file in seen and seen[0] == cur_file:
            break   # second profiled launch: same layout again
        seen.append(cur_file); continue
    if r and r[0] == "Line No":
        hdr = r; iex = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples"); continue
    if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
        continue
    key = (cur_file, int(r[0]))
    try:
        agg[key][0] += int(r[iex] or 0); agg[key][1] += int(r[isamp] or 0)
    except ValueError:
        pass
    agg[key][2] = r[1].strip()[:100]
tot_i = sum(v[0] for v in agg.values()); tot_s = sum(v[1] for v in agg.values())
print(f"total inst {tot_i}  samples {tot_s}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k[0]:16s}:{k[1]:4d} inst {v[0]:9d} {100*v[0]/max(tot_i,1):5.1f}%  samp {100*v[1]/max(tot_s,1):5.1f}%  {v[2]}")
