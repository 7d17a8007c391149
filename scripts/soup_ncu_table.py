#!/usr/bin/env python
"""Per-structure DRAM / L2 figures of the triangle-soup sweep (BASELINE config 5) from the ncu pass
  CONFIG5_REPS=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum
                     --clock-control none -k regex:k_trace --csv --log-file soup_ncu.csv python scripts/config5_soup_sweep.py 6 22 6
joined with the records the same run wrote (order of the timed launches = order of the records).
usage: scripts/soup_ncu_table.py soup_ncu.csv run.json [out.json [plain_sweep.json]]
(the Mrays/s of a run under ncu mean nothing: they are taken from the plain sweep's record of the same n and structure)"""
import csv
import json
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if r]
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
launches = {}
for r in rows[hi + 1:]:
    if len(r) != len(hdr):
        continue
    d = launches.setdefault(int(r[col["ID"]]), {"kernel": r[col["Kernel Name"]]})
    v = float(r[col["Metric Value"]].replace(",", ""))
    u = r[col["Metric Unit"]].lower()
    scale = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "nsecond": 1e-9,
             "s": 1, "second": 1, "sector": 1}.get(u, 1)
    d[r[col["Metric Name"]]] = v * scale
# the timed launches are the ones without test counting: k_trace<MODE, COUNT = 0> / k_trace_bvh<COUNT = 0, EARLY>
import re


def is_timed(kernel):
    m = re.search(r"k_trace(_bvh)?<\(?(?:int|bool)?\)?\s*(\w+),\s*\(?(?:int|bool)?\)?\s*(\w+)>", kernel)
    if not m:
        return False
    count = m.group(2) if m.group(1) else m.group(3)
    return count in ("0", "false")


timed = [l for _, l in sorted(launches.items()) if is_timed(l["kernel"])]
recs = [r for r in json.load(open(sys.argv[2])) if "ms" in r]
plain = {}
if len(sys.argv) > 4:
    plain = {(r["n"], r["structure"]): r for r in json.load(open(sys.argv[4])) if "ms" in r}
out = []
for rec, l in zip(recs, timed):
    nrays = rec["mrays_s"] * rec["ms"] * 1e3
    rec = dict(rec, mrays_s=plain.get((rec["n"], rec["structure"]), {}).get("mrays_s", float("nan")), nrays=nrays)
    secs = l["gpu__time_duration.sum"]
    dram = l["dram__bytes_read.sum"] + l["dram__bytes_write.sum"]
    l2 = l["lts__t_sectors.sum"] * 32
    out.append({"n": rec["n"], "structure": rec["structure"], "kernel": l["kernel"][:60], "launch_ms_under_ncu": secs * 1e3, "mrays_s_plain_sweep": rec["mrays_s"],
                "dram_gb_s": dram / secs / 1e9, "l2_gb_s": l2 / secs / 1e9, "dram_bytes_per_ray": dram / rec["nrays"], "rays_in_launch": rec["nrays"],
                "algorithmic_bytes_per_ray": rec["algorithmic_bytes_per_ray"], "as_tests_per_ray": rec["as_tests_per_ray"], "tri_tests_per_ray": rec["tri_tests_per_ray"]})
    print(f"n {rec['n']:>9}  {rec['structure']:26s} {rec['mrays_s']:8.1f} Mrays/s  DRAM {out[-1]['dram_gb_s']:7.1f} GB/s  L2 {out[-1]['l2_gb_s']:7.1f} GB/s  "
          f"DRAM B/ray {out[-1]['dram_bytes_per_ray']:8.1f}  algorithmic B/ray {rec['algorithmic_bytes_per_ray']:8.1f}")
if len(timed) != len(recs):
    print(f"warning: {len(timed)} timed launches vs {len(recs)} records", file=sys.stderr)
if len(sys.argv) > 3:
    json.dump(out, open(sys.argv[3], "w"), indent=1)
