#!/usr/bin/env python
"""Folds one ncu capture of the path kernel into profiles/r2_ncu_counters.json — the measured-bound figures bench.py puts
into `roofline` (lanes per issued instruction, issue-slot utilisation, DRAM and L2 GB/s, DRAM bytes per ray).
usage: scripts/ncu_counters.py <key: lucy_bvh|lucy_subgrid|lucy_subbeams|cornell_bvh...> <report.ncu-rep> <rays in the captured launch> "<command captured>"
Read here, no GPU needed.  The capture is `ncu --set full --clock-control none -k regex:k_mega -s <warm-up launches> -c 1`."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
key, rep, rays, cmd = sys.argv[1], sys.argv[2], float(sys.argv[3]), sys.argv[4]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, r = rows[0], rows[1], rows[2]


def val(name):
    i = hdr.index(name)
    v = float(r[i].replace(",", ""))
    u = units[i].lower()
    scale = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3,
             "msecond": 1e-3, "s": 1.0, "second": 1.0, "nsecond": 1e-9}.get(u, 1.0)
    return v * scale


secs = val("gpu__time_duration.sum")
dram = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
l2 = val("lts__t_sectors.sum") * 32.0  # L2 sectors are 32 bytes
out = {
    "kernel": r[hdr.index("Kernel Name")], "source": f"profiles/{os.path.basename(rep).replace('.ncu-rep', '.txt')} ({cmd})",
    "launch_ms_under_ncu": secs * 1e3, "rays_in_launch": rays,
    "lanes_per_inst": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "warps_active_pct": val("sm__warps_active.avg.pct_of_peak_sustained_active"),
    "registers": val("launch__registers_per_thread"),
    "dram_bytes": dram, "dram_bytes_read": val("dram__bytes_read.sum"), "dram_bytes_write": val("dram__bytes_write.sum"),
    "dram_gbs": dram / secs / 1e9, "dram_bytes_per_ray": dram / rays,
    "l2_bytes": l2, "l2_gbs": l2 / secs / 1e9,
    "l1_hit_pct": val("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": val("lts__t_sector_hit_rate.pct"),
    "dram_pct_of_peak": val("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    "inst_executed": val("smsp__inst_executed.sum"),
}
path = os.path.join(ROOT, "profiles", "r2_ncu_counters.json")
tab = json.load(open(path)) if os.path.isfile(path) else {}
tab[key] = out
json.dump(tab, open(path, "w"), indent=1, sort_keys=True)
print(key, json.dumps(out, indent=1))
