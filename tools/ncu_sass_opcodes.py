#!/usr/bin/env python
"""Executed warp-instructions per SASS opcode from an `ncu --page source --csv --print-source cuda,sass` dump."""
import csv, sys, re
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
cnt = Counter(); hdr = None; seen = []; seen_addr = set()
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        f = r[1]
        if seen and seen[0] == f: break
        seen.append(f); continue
    if r and r[0] == "Line No":
        hdr = r; iex = hdr.index("Instructions Executed"); continue
    if hdr is None or len(r) < len(hdr) or r[0] != "" or not r[2].startswith("0x"): continue
    if r[2] in seen_addr: continue
    seen_addr.add(r[2])
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[3])
    if not m: continue
    op = m.group(2)
    try: cnt[op] += int(r[iex])
    except ValueError: pass
tot = sum(cnt.values())
print("total", tot)
for op, c in cnt.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 40):
    print(f"{op:28s} {c:10d} {100*c/tot:5.1f}%")
