#!/usr/bin/env python
"""DRAM traffic of the frame kernel over CONSECUTIVE launches (canvas write-back inside the window).

    python tools/traffic.py run  <workload>     (GPU box) -> gpurun_out/traffic_<workload>.csv
    python tools/traffic.py fold <workload>...  (anywhere) -> profiles/dram_traffic.json

`run` profiles 16 back-to-back launches of the bench's device-resident ring with `ncu --cache-control none`
(caches are not flushed between launches, so what one frame leaves dirty in L2 is written back while the next
frames run and lands in their counters); the per-launch average of dram__bytes_read + dram__bytes_write is what
bench.py prints as roofline.traffic."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(w):
    cmd = ["python", "bench.py", "--steps", "2", "--warmup", "3", "--frames-per-step", "16", "--no-e2e",
           "--no-cpu-baseline", "--workload", w]
    subprocess.check_call(cmd, cwd=ROOT, stdout=subprocess.DEVNULL)          # plain run first
    out = os.path.join(ROOT, "gpurun_out", f"traffic_{w}.csv")
    subprocess.check_call(["ncu", "--metrics", "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum",
                           "--cache-control", "none", "--clock-control", "none", "-k", "regex:stitch_fast", "-s", "32",
                           "-c", "16", "--csv", "--log-file", out] + cmd, cwd=ROOT, stdout=subprocess.DEVNULL)
    print(out)


def fold(ws):
    res = {"how": "ncu --cache-control none --clock-control none, 16 consecutive launches of the frame kernel in the "
                  "bench's device-resident ring (8 frame sets), average per launch of dram__bytes_read.sum + "
                  "dram__bytes_write.sum (tools/traffic.py)"}
    for w in ws:
        path = os.path.join(ROOT, "gpurun_out", f"traffic_{w}.csv")
        rows = [r for r in csv.reader(open(path)) if len(r) > 5]
        hdr = rows[0]
        iname, ival, iunit, iid = hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit"), hdr.index("ID")
        per = {}
        for r in rows[1:]:
            v = float(r[ival].replace(",", ""))
            u = r[iunit]
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            per.setdefault(r[iid], {})[r[iname]] = v * scale
        rd = [p["dram__bytes_read.sum"] for p in per.values()]
        wr = [p["dram__bytes_write.sum"] for p in per.values()]
        res[w] = int((sum(rd) + sum(wr)) / len(rd))
        res[w + "_detail"] = {"launches": len(rd), "read_avg": int(sum(rd) / len(rd)), "write_avg": int(sum(wr) / len(wr))}
    json.dump(res, open(os.path.join(ROOT, "profiles", "dram_traffic.json"), "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    if sys.argv[1] == "run":
        run(sys.argv[2])
    else:
        fold(sys.argv[2:])
