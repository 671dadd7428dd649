#!/usr/bin/env python
"""Executed warp instructions and stall samples of one kernel PER SOURCE LINE: joins the per-SASS-instruction counters of
an ncu report (`--page source --csv`) with the line table of the object file (`nvdisasm -g`, needs -lineinfo).
usage: scripts/ncu_by_line.py <rep> <kernel regex> <object.o> [topN] [source-file-substring]
"""
import collections, csv, io, re, subprocess, sys

rep, kre, obj = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
only = sys.argv[5] if len(sys.argv) > 5 else None

out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kre}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
iex, ismp, isrc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
body = [r for r in rows[hi + 1:] if len(r) > iex and r[iex].isdigit()]
half = len(body) // 2
if half and body[0][0] == body[half][0]:     # the kernel matched twice (two launches): keep the first
    body = body[:half]
base = int(body[0][0], 16)
per_addr = {int(r[0], 16) - base: (int(r[iex]), int(r[ismp] or 0), r[isrc]) for r in body}

# nvdisasm -g: "//## File "path", line N" comments precede the instructions they belong to; functions start at ".text.<name>:"
# a host object holds the cubin as a fat-binary member: extract it first
import glob, os, tempfile
cubin = obj
if not obj.endswith(".cubin"):
    tmp = tempfile.mkdtemp(prefix="ncu_by_line_")
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
    found = glob.glob(os.path.join(tmp, "*.cubin"))
    assert found, f"no cubin inside {obj}"
    cubin = found[0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
line_of, cur, infn = {}, None, False
fn_re = re.compile(r"^\s*\.text\.(\S+):")
for l in dis.splitlines():
    m = fn_re.match(l)
    if m:
        infn = re.search(kre, m.group(1)) is not None and not line_of
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", l)
    if m and cur:
        line_of[int(m.group(1), 16)] = cur

agg = collections.defaultdict(lambda: [0, 0, 0])
tot_ex = tot_smp = 0
for a, (ex, smp, _src) in per_addr.items():
    key = line_of.get(a, ("?", 0))
    agg[key][0] += ex; agg[key][1] += smp; agg[key][2] += 1
    tot_ex += ex; tot_smp += smp
src_cache = {}
def text(f, n):
    if f not in src_cache:
        try:
            src_cache[f] = open(f"3dgrut_b200/csrc/{f}").read().splitlines()
        except OSError:
            src_cache[f] = []
    s = src_cache[f]
    return s[n - 1].strip()[:90] if 0 < n <= len(s) else ""
print(f"warp instructions {tot_ex}, samples {tot_smp}, SASS instructions {len(per_addr)}, mapped {sum(1 for a in per_addr if a in line_of)}")
print(f"{'file:line':28s} {'sass':>5s} {'executed%':>9s} {'samples%':>9s}")
for (f, n), (ex, smp, cnt) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    if only and only not in f:
        continue
    print(f"{f + ':' + str(n):28s} {cnt:5d} {100 * ex / max(tot_ex, 1):8.2f}% {100 * smp / max(tot_smp, 1):8.2f}%  {text(f, n)}")
