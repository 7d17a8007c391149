#!/usr/bin/env python
"""Per-region view of an .ncu-rep's SASS page (read here, no GPU): contiguous address ranges with their share of the issued
warp instructions, their active lanes per instruction and their stall samples.  Regions are cut where the executed-instruction
count of neighbouring instructions changes by more than 25 % (basic-block boundaries show up as count steps); the source
line most samples of a region map to is printed beside it (needs -lineinfo).
usage: scripts/ncu_sass_regions.py report.ncu-rep [min_share_pct]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
minshare = float(sys.argv[2]) if len(sys.argv) > 2 else 0.7
run = lambda *a: subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-count", "1", *a], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(run("--print-source", "sass"))))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
idx = {h: i for i, h in enumerate(hdr)}
ins = []
for r in rows[hi + 1:]:
    if len(r) != len(hdr) or r[0] == "Address":
        continue
    g = lambda k: int(r[idx[k]] or 0)
    ins.append((r[0], r[idx["Source"]].strip(), g("Instructions Executed"), g("Thread Instructions Executed"), g("# Samples"), g("stall_long_sb")))
tot = sum(i[2] for i in ins)
tots = sum(i[4] for i in ins)
print(f"{len(ins)} SASS instructions, {tot} warp instructions, {sum(i[3] for i in ins) / tot:.2f} lanes each, {tots} samples")
reg, cur = [], [0]
for k in range(1, len(ins)):
    a, b = ins[k - 1][2], ins[k][2]
    if max(a, b) > 1.25 * max(1, min(a, b)):
        reg.append(cur)
        cur = []
    cur.append(k)
reg.append(cur)
for g in reg:
    n = sum(ins[k][2] for k in g)
    if n / tot * 100 < minshare:
        continue
    th = sum(ins[k][3] for k in g)
    sm = sum(ins[k][4] for k in g)
    lsb = sum(ins[k][5] for k in g)
    ops = collections.Counter()
    for k in g:
        op = ins[k][1].split()
        ops[(op[1] if op[0].startswith("@") else op[0]).split(".")[0]] += 1
    top = " ".join(f"{o}:{c}" for o, c in ops.most_common(6))
    print(f"[{g[0]:5d}-{g[-1]:5d}] {len(g):4d} instr  share {n / tot * 100:5.1f}%  lanes {th / max(1, n):5.1f}  samples {sm / max(1, tots) * 100:5.1f}% "
          f"(long_sb {lsb / max(1, sm) * 100:3.0f}%)  exec/instr {n / len(g):.3g}  {top}")
