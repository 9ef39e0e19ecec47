#!/bin/bash
# usage (on the GPU box): tools/variants.sh <tag> <lib name>...   -- short device-resident bench per library build
# (image-stitching_b200/lib_<name>.so built with other macros; `default` = libb200stitch.so)
T=$1; shift
for v in "$@"; do
  L=image-stitching_b200/lib_$v.so; [ "$v" = default ] && L=image-stitching_b200/libb200stitch.so
  B200_STITCH_LIB=$PWD/$L python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline ${BENCH_ARGS} > gpurun_out/${T}_$v.json 2> gpurun_out/${T}_$v.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${T}_$v.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("$v", "pipelined us", round(r["kernel_us_pipelined"],2), "isolated", round(r["kernel_us_isolated"],2), "frac", round(r["frac"],3), d["parity"], d["run"]["kernel"])
except Exception as e:
    print("$v failed", e); print(open("gpurun_out/${T}_$v.err").read()[-800:])
PY
done
