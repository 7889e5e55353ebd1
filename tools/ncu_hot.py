#!/usr/bin/env python
"""Top stall sites of an ncu source page: python tools/ncu_hot.py src.csv [N] [section]

A report with several launches holds one table per launch and view (SASS, then source); ``section`` picks one
(default 0)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
sec = int(sys.argv[3]) if len(sys.argv) > 3 else 0
heads = [i for i, r in enumerate(rows) if r and r[0] in ("Address", "#")]
hi = heads[sec]
end = heads[sec + 1] if sec + 1 < len(heads) else len(rows)
h = rows[hi]
ix = {k: i for i, k in enumerate(h)}
body = [r for r in rows[hi + 1:end] if len(r) == len(h)]
tot = sum(int(r[ix["# Samples"]] or 0) for r in body)
print("sections", len(heads), "total samples", tot)
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
for pos, r in sorted(enumerate(body), key=lambda t: -int(t[1][ix["# Samples"]] or 0))[:n]:
    s = int(r[ix["# Samples"]] or 0)
    top = sorted(((int(r[ix[k]] or 0), k) for k in stalls), reverse=True)[:2]
    print(f"{pos:5d} {100*s/tot:5.1f}% {r[ix['Source']][:70]:70s} {top[0][1]}={top[0][0]} {top[1][1]}={top[1][0]}")
