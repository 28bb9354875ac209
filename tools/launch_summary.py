#!/usr/bin/env python
"""Condense an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel totals (profiles/)."""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ik, iv, iu, ib, ig = (hdr.index(k) for k in ("Kernel Name", "Metric Value", "Metric Unit", "Block Size", "Grid Size"))
tot, cnt, geo = collections.Counter(), collections.Counter(), {}
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ik]).replace("void ", "").replace("vdp::", "")
    v = float(r[iv].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0}.get(r[iu], 1e-6)
    name = (name, r[ig], r[ib])      # the same kernel appears with several geometries (resident batch / chunks)
    tot[name] += v
    cnt[name] += 1
total = sum(tot.values())
w = csv.writer(sys.stdout)
w.writerow(["kernel", "launches", "ms_total", "ms_per_launch", "share_pct", "grid", "block"])
for k, v in tot.most_common():
    w.writerow([k[0], cnt[k], "%.4f" % v, "%.4f" % (v / cnt[k]), "%.2f" % (100 * v / total), k[1], k[2]])
w.writerow(["TOTAL", sum(cnt.values()), "%.4f" % total, "", "100", "", ""])
