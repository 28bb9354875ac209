#!/usr/bin/env python
"""Warp-stall sample totals by reason for one kernel from `ncu --page source --csv` (+ top stalled instructions)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(r for r in rows if "Instructions Executed" in r)
cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
isrc, ins = hdr.index("Source"), hdr.index("# Samples")
tot = {h: 0 for _, h in cols}
top = []
for r in rows:
    if len(r) < len(hdr) or not r[ins].isdigit():
        continue
    for i, h in cols:
        tot[h] += int(r[i] or 0)
    top.append((int(r[ins]), r[isrc].strip(), {h: int(r[i] or 0) for i, h in cols if int(r[i] or 0) > 0}))
s = sum(tot.values())
print("samples", s)
for h, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    if v:
        print(f"  {h:24s} {100 * v / s:5.1f}%")
top.sort(key=lambda t: -t[0])
for n, src, d in top[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    print(f"{100 * n / s:5.1f}%  {src[:60]:60s} {d}")
