#!/usr/bin/env python
"""usage: scripts/ncu_stalls.py rep kernel_regex [topN]   -- stall-reason mix and hottest SASS lines of one kernel"""
import csv, io, subprocess, sys
rep, k = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{k}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
iex, ismp, isrc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[hi + 1:] if len(r) > max(cols) and r[iex].isdigit()]
half = len(body) // 2
if half and body[0][0] == body[half][0]:
    body = body[:half]
tots = sum(int(r[ismp] or 0) for r in body)
tot = {hdr[i]: 0 for i in cols}
for r in body:
    for i in cols:
        tot[hdr[i]] += int(r[i] or 0)
s = sum(tot.values())
print("warp instructions:", sum(int(r[iex]) for r in body), " samples:", tots)
print(", ".join(f"{kk[6:]} {100*v/s:.1f}%" for kk, v in sorted(tot.items(), key=lambda kv: -kv[1])[:9]))
for r in sorted(body, key=lambda r: -int(r[ismp] or 0))[:top]:
    why = max(cols, key=lambda i: int(r[i] or 0))
    print(f"{100*int(r[ismp])/max(tots,1):5.1f}% ex={r[iex]:>10s} {hdr[why][6:]:14s} {r[0][-5:]} {r[isrc][:80]}")
