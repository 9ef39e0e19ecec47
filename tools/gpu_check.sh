#!/bin/bash
# usage (on the GPU box): tools/gpu_check.sh <tag>   -- parity tests, short bench of both staged kernels, one ncu capture
T=$1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${T}_bench_v2.json 2> gpurun_out/${T}_bench_v2.err
BS_DEBUG_NO_COMPACT=1 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${T}_bench_v1.json 2> gpurun_out/${T}_bench_v1.err
python - <<PY
import json
for v in ("v2","v1"):
    try:
        d=json.loads(open("gpurun_out/${T}_bench_%s.json"%v).read().strip().splitlines()[-1])
        r=d["roofline"]; print(v, d["run"]["kernel"], "pipelined us", round(r["kernel_us_pipelined"],2), "isolated", round(r["kernel_us_isolated"],2), "frac", round(r["frac"],3), d["parity"])
    except Exception as e: print(v, "failed", e); print(open("gpurun_out/${T}_bench_%s.err"%v).read()[-1500:])
PY
tools/prof.sh $T
