#!/usr/bin/env python
"""FP32 work a kernel EXECUTED, from the per-instruction counters of an ncu report (`--page source`): thread-level executed
counts of every FP32 arithmetic SASS instruction, packed (FFMA2 / FMUL2 / FADD2: two results per thread) and scalar, turned
into FLOP (an FMA counts 2).  ncu's op_{fadd,fmul,ffma}_pred_on counters see only the scalar forms.
usage: scripts/ncu_flops.py <rep> <kernel regex>"""
import collections, csv, io, subprocess, sys

rep, kre = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kre}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# one section per matching launch, each introduced by its own header row: keep the launch that executed most
sections, hdr = [], None
for r in rows:
    if r and r[0] == "Address":
        hdr = r
        sections.append([])
    elif hdr is not None and sections:
        sections[-1].append(r)
iex, ith, isrc = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("Source")
sections = [[r for r in sec if len(r) > iex and r[iex].isdigit()] for sec in sections]
body = max(sections, key=lambda sec: sum(int(r[iex]) for r in sec))
FLOP = {"FFMA2": 4, "FMUL2": 2, "FADD2": 2, "FFMA": 2, "FMUL": 1, "FADD": 1}
warp = collections.Counter(); thread = collections.Counter()
tot_warp = 0
for r in body:
    src = r[isrc].strip()
    while src.startswith("@"):
        src = src.split(None, 1)[1]
    op = src.split()[0].split(".")[0]
    tot_warp += int(r[iex])
    if op in FLOP:
        warp[op] += int(r[iex]); thread[op] += int(r[ith])
flop = sum(thread[o] * FLOP[o] for o in thread)
print(f"warp instructions executed: {tot_warp}")
for o in FLOP:
    print(f"  {o:6s} warp-level {warp[o]:>12d}  thread-level {thread[o]:>14d}  FLOP {thread[o] * FLOP[o]:>16d}")
fp_warp = sum(warp.values())
print(f"FP32 arithmetic: {fp_warp} warp instructions ({100 * fp_warp / tot_warp:.1f} % of all), {flop:.4e} FLOP executed; "
      f"FMA-pipe cycles per SM sub-partition if packed = 2, scalar = 1: {2 * (warp['FFMA2'] + warp['FMUL2'] + warp['FADD2']) + warp['FFMA'] + warp['FMUL'] + warp['FADD']}")
