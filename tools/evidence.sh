#!/bin/bash
# usage (on the GPU box): tools/evidence.sh <tag>
# Everything the round-end driver runs, plus the evidence kept under profiles/ (copied there by hand afterwards):
# bench lines of every BASELINE.json config, the lens-model line, the reference arm, the ncu launch list, one full
# ncu capture of the frame kernel and of the lens kernel, DRAM traffic over consecutive launches.
T=$1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -3 > gpurun_out/${T}_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
for w in cfg2 cfg5 cfg1; do
  python bench.py --workload $w --steps 20 --cpu-seconds 6 > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err
done
python bench.py --lens --steps 20 --cpu-seconds 8 > gpurun_out/${T}_bench_lens.json 2> gpurun_out/${T}_bench_lens.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${T}_ncu_list.log 2>&1
tools/prof.sh $T
for w in cfg3 cfg2 cfg5; do python tools/traffic.py run $w > /dev/null 2>&1; done
tools/prof_lens.sh $T
python tools/feather_bench.py cfg3 > gpurun_out/${T}_feather_bench.json 2>/dev/null; python tools/feather_bench.py cfg2 >> gpurun_out/${T}_feather_bench.json 2>/dev/null
python tools/copy_patterns.py > gpurun_out/${T}_copy_patterns.json 2>/dev/null
tools/jpeg_launches.sh $T
tail -2 gpurun_out/${T}_pytest.log; tail -1 gpurun_out/${T}_smoke.log
python - <<PY
import json
for n in ("bench","bench_cfg2","bench_cfg5","bench_cfg1","bench_lens","bench_reference"):
    try:
        d=json.loads(open("gpurun_out/${T}_%s.json"%n).read().strip().splitlines()[-1])
        r=d.get("roofline") or {}
        print(n, "value", round(d["value"],1), "e2e", round((d.get("e2e") or {}).get("value",0),1), "us", round(r.get("kernel_us_pipelined",0),2), "iso", round(r.get("kernel_us_isolated",0),2), "frac", round(r.get("frac",0),3), d.get("parity"), (d.get("undistort") or {}).get("us_per_camera_frame"))
    except Exception as e: print(n, "failed", e)
PY
