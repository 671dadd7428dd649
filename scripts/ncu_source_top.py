#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export: SASS instructions sorted by address with executed counts.
usage: ncu -i rep --page source --csv --kernel-name regex:K > src.csv ; scripts/ncu_source_top.py src.csv [min_share]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
# find header row
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ia, isrc, iex, ith, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
body = [r for r in rows[hi + 1:] if len(r) > iex and r[iex].isdigit()]
tot = sum(int(r[iex]) for r in body)
tots = sum(int(r[ismp] or 0) for r in body)
print(f"total warp instructions {tot}, samples {tots}, sass lines {len(body)}")
for r in body:
    ex = int(r[iex]); sm = int(r[ismp] or 0)
    thr = int(r[ith]) / ex if ex else 0
    print(f"{r[ia][-5:]} ex={ex:>11d} {100*ex/tot:5.2f}% smp={100*sm/max(tots,1):5.2f}% thr={thr:4.1f}  {r[isrc][:110]}")
