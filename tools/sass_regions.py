#!/usr/bin/env python
"""SASS instruction count of one kernel by outermost source line range (nvdisasm -gi dump).
usage: sass_regions.py dump.txt kernel_symbol_substring file.cu bucket_size"""
import re, sys
from collections import Counter
lines = open(sys.argv[1]).read().split('\n')
sym = sys.argv[2]; src = sys.argv[3]; bucket = int(sys.argv[4]) if len(sys.argv) > 4 else 20
start = None
for i, l in enumerate(lines):
    if l.startswith('.text.') and l.endswith(':') and sym in l: start = i; continue
    if start is not None and l.startswith('\t.section'): end = i; break
else: end = len(lines)
cnt = Counter(); cur = None; pending = []
tot = 0
for l in lines[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        f, n, rest = m.group(1), int(m.group(2)), m.group(3)
        # with -gi, inlined code has extra lines "inlined at File..., line N"; keep outermost in src
        if 'inlined at' in l:
            m2 = re.findall(r'inlined at "([^"]+)", line (\d+)', l)
            for ff, nn in m2:
                if ff.endswith(src): n = int(nn); f = ff
        if f.endswith(src): cur = n
        continue
    m = re.search(r'//## +inlined at "([^"]+)", line (\d+)', l)
    if m:
        if m.group(1).endswith(src): cur = int(m.group(2))
        continue
    if re.match(r'\s+/\*[0-9a-f]{4,5}\*/', l):
        tot += 1; cnt[(cur or 0) // bucket * bucket] += 1
print('total', tot)
for k in sorted(cnt): print(f"lines {k:4d}-{k+bucket-1:4d}: {cnt[k]}")
