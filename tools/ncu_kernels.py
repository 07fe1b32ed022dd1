#!/usr/bin/env python
"""Per-kernel table of an `ncu --set full` report: duration, DRAM bytes, occupancy, issue-slot use, registers, DRAM and
tensor-pipe utilisation, and the two leading warp-stall reasons.  Usage: python tools/ncu_kernels.py REPORT.ncu-rep
[--traffic OUT.json KERNEL_SUBSTRING]   (the optional part writes the DRAM bytes of one launch for bench.py)"""
import csv
import json
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
ci = {h: i for i, h in enumerate(hdr)}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def val(r, name, conv=False):
    if name not in ci or not r[ci[name]]:
        return float("nan")
    v = float(r[ci[name]].replace(",", ""))
    return v * scale.get(units[ci[name]], 1.0) if conv else v


stalls = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")
          and "not_issued" not in h]
print(f"{'kernel':44s} {'ms':>8s} {'rd GB':>7s} {'wr GB':>7s} {'occ%':>5s} {'issue%':>6s} {'regs':>4s} {'dram%':>5s} {'tens%':>5s}  leading stalls")
for r in rows[2:]:
    name = r[ci["Kernel Name"]].replace("void ", "").replace("cip::<unnamed>::", "").replace("cip::", "")[:44]
    st = sorted(((val(r, s), s.split("stalled_")[1].split("_per_")[0]) for s in stalls), reverse=True)
    st = [s for s in st if s[1] != "selected"][:2]
    print(f"{name:44s} {val(r, 'gpu__time_duration.sum', True):8.3f} {val(r, 'dram__bytes_read.sum', True) / 1e9:7.3f} "
          f"{val(r, 'dram__bytes_write.sum', True) / 1e9:7.3f} {val(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):5.1f} "
          f"{val(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):6.1f} {val(r, 'launch__registers_per_thread'):4.0f} "
          f"{val(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):5.1f} "
          f"{val(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'):5.1f}  "
          + ", ".join(f"{n} {v:.2f}" for v, n in st))
if len(sys.argv) > 4 and sys.argv[2] == "--traffic":
    for r in rows[2:]:
        if sys.argv[4] in r[ci["Kernel Name"]]:
            d = {"kernel": r[ci["Kernel Name"]], "source": rep,
                 "dram_bytes": int(val(r, "dram__bytes_read.sum", True) + val(r, "dram__bytes_write.sum", True))}
            open(sys.argv[3], "w").write(json.dumps(d) + "\n")
            break
