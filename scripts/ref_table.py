#!/usr/bin/env python
"""Tabulates gpurun_out/parity_ref.json (written by scripts/ref_report.py) as text: one row per statistic, one column per case."""
import json
import sys

path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/parity_ref.json"
d = json.load(open(path))
keys = []
for c, s in d.items():
    for k in s:
        if k not in keys and k not in ("case", "trace"):
            keys.append(k)
cols = list(d)
for i, c in enumerate(cols):
    print(f"[{i}] {c}")
print(f"{'statistic':55s}", *[f"[{i}]".rjust(9) for i in range(len(cols))])
for k in keys:
    cells = []
    for c in cols:
        v = d[c].get(k)
        cells.append((f"{v:.2e}" if isinstance(v, float) else str(v))[:9].rjust(9))
    print(f"{k:55s}", *cells)
