#!/usr/bin/env python
"""Compact form of an `ncu --metrics gpu__time_duration.sum --csv` launch list: one row per launch of this
library's kernels (in launch order), one summary row per other kernel name (torch's frame generators etc.).
usage: python tools/launch_summary.py gpurun_out/<tag>_launches.csv > profiles/<name>_launches.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10]
hdr = rows[0]
iid, iname, ival, iunit = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
ours = ("stitch_", "plan_kernel", "undistort_", "lens_plan")
other = collections.OrderedDict()
w = csv.writer(sys.stdout)
w.writerow(["launch_id", "kernel", "gpu__time_duration_us"])
n_ours = 0
tot_ours = 0.0
for r in rows[1:]:
    name = r[iname]
    us = float(r[ival].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[iunit], 1.0)
    if any(k in name for k in ours):
        short = name.split("(")[0].split("::")[-1]
        w.writerow([r[iid], short, f"{us:.3f}"])
        n_ours += 1
        tot_ours += us
    else:
        k = name[:70]
        c = other.setdefault(k, [0, 0.0, r[iid]])
        c[0] += 1
        c[1] += us
w.writerow([])
w.writerow(["# other kernels (not in any timed region: torch builds the synthetic frames before it)", "launches", "total_us", "first_launch_id"])
for k, (n, us, first) in other.items():
    w.writerow([k, n, f"{us:.1f}", first])
w.writerow([])
w.writerow([f"# this library: {n_ours} launches, {tot_ours:.1f} us in total (cold-cache, serialised by the profiler)"])
