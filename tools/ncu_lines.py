#!/usr/bin/env python
"""Aggregate warp-stall samples per CUDA source line from
`ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > file.csv`."""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, hdr = None, None
agg = defaultdict(lambda: [0, 0, ""])
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) > 6 and r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) == len(hdr) and r[2] == "-":   # a CUDA source line row (aggregated)
        try:
            s = int(r[6]); ie = int(r[7])
        except ValueError:
            continue
        a = agg[(cur, r[0])]
        a[0] += s; a[1] += ie; a[2] = r[1][:100]
tot = sum(a[0] for a in agg.values())
print("total samples", tot)
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{a[0]:6d} {100.0 * a[0] / max(tot, 1):5.1f}%  inst={a[1]:8d}  {f}:{ln}  {a[2]}")
