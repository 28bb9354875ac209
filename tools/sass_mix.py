#!/usr/bin/env python
"""Opcode mix (executed warp instructions, stall samples) of one kernel from `ncu --page source --csv`."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(r for r in rows if "Instructions Executed" in r)
ia, isrc, ist = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
tot = 0
byop, samp = collections.Counter(), collections.Counter()
for r in rows:
    if len(r) <= ia or not r[ia].isdigit():
        continue
    n = int(r[ia])
    tot += n
    op = r[isrc].split()
    o = op[0] if not op[0].startswith("@") else op[1]
    o = o.split(".")[0]
    byop[o] += n
    samp[o] += int(r[ist])
print("total warp instructions", tot)
st = sum(samp.values())
for o, n in byop.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 22):
    print(f"{o:10s} {n:12d} {100 * n / tot:5.1f}%   samples {100 * samp[o] / max(st, 1):5.1f}%")
