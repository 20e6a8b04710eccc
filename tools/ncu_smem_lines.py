#!/usr/bin/env python
"""Shared-memory wavefronts per CUDA source line (actual / ideal) from
`ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > file.csv`."""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
cur, hdr = None, None
agg = defaultdict(lambda: [0, 0, ""])
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) > 6 and r[0] == "Line No":
        hdr = r
        iw, ii = hdr.index("L1 Wavefronts Shared"), hdr.index("L1 Wavefronts Shared Ideal")
        continue
    if hdr and len(r) == len(hdr) and r[2] == "-":
        try:
            w = int(r[iw]); i = int(r[ii])
        except ValueError:
            continue
        a = agg[(cur, r[0])]
        a[0] += w; a[1] += i; a[2] = r[1][:100]
tot = sum(a[0] for a in agg.values())
print("total shared wavefronts", tot, "ideal", sum(a[1] for a in agg.values()))
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{a[0]:10d} {100.0 * a[0] / max(tot, 1):5.1f}%  ideal={a[1]:10d}  {f}:{ln}  {a[2]}")
