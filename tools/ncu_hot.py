#!/usr/bin/env python
"""Opcode mix per ray-step of the hot region of an ncu source-page CSV.
usage: ncu -i rep --page source --csv > src.csv; python tools/ncu_hot.py src.csv <ray_steps_per_launch>"""
import csv, re, sys
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
steps = float(sys.argv[2]) if len(sys.argv) > 2 else 5.11e9
hdr, data = rows[1], rows[2:]
isrc, iex = hdr.index('Source'), hdr.index('Instructions Executed')
W = steps / 32
tot = sum(float(r[iex]) for r in data)
hot = [r for r in data if float(r[iex]) > 0.5 * W]
ops = Counter()
for r in hot:
    t = re.sub(r'^@!?U?P\d+\s+', '', r[isrc])
    ops[t.split()[0].split('.')[0]] += float(r[iex]) / W
print("warp-instructions per ray-step: total %.1f, hot region %.1f (%d SASS lines)" % (tot / W, sum(ops.values()), len(hot)))
print({k: round(v, 1) for k, v in ops.most_common(30)})
cols = [c for c in hdr if c.startswith('stall_') and 'Not Issued' not in c]
tots = {c: sum(float(r[hdr.index(c)] or 0) for r in data) for c in cols}
s = sum(tots.values())
print("stall samples:", {k: round(v / s, 3) for k, v in sorted(tots.items(), key=lambda x: -x[1])[:9]})
