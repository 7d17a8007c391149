#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key raw metrics per launch, stall mix and instruction mix.
usage: scripts/ncu_summary.py report.ncu-rep [out.txt]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
out = open(sys.argv[2], "w") if len(sys.argv) > 2 else sys.stdout
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "lts__t_bytes.sum"]
for r in rows[2:]:
    print("---- launch", file=out)
    for k in want:
        for i, h in enumerate(hdr):
            if h == k:
                print(f"{k} = {r[i][:110]} {units[i]}", file=out)
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--launch-count", "1"],
                      stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(sass)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
idx = {h: i for i, h in enumerate(hdr)}
first, seen = [], set()
for r in rows[hi + 1:]:
    if len(r) != len(hdr) or r[0] == "Address":
        continue
    if r[0] in seen:
        break
    seen.add(r[0])
    first.append(r)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter()
mix, act = collections.Counter(), collections.Counter()
for r in first:
    for s in stalls:
        try:
            tot[s] += int(r[idx[s]])
        except ValueError:
            pass
    src = r[idx["Source"]].split()
    op = (src[1] if src[0].startswith("@") else src[0]).split(".")[0]
    n = int(r[idx["Instructions Executed"]] or 0)
    mix[op] += n
    act[op] += int(r[idx["Thread Instructions Executed"]] or 0)
T = sum(tot.values())
print("---- stall samples (first launch)", file=out)
for s, v in tot.most_common(8):
    print(f"{s} {v} {v / max(1, T) * 100:.1f}%", file=out)
ti = sum(mix.values())
print(f"---- instruction mix (first launch): {ti} warp instructions, {sum(act.values()) / max(1, ti):.2f} active lanes on average", file=out)
for k, v in mix.most_common(16):
    print(f"{k:10s} {v:11d} {v / ti * 100:5.1f}%  lanes {act[k] / max(1, v):.1f}", file=out)
