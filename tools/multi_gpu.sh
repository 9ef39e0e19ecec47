#!/bin/bash
# usage (on a multi-GPU box): tools/multi_gpu.sh <tag> <N>...   -- bench lines at N GPUs (cfg3, and cfg5 at the largest N), pool test
T=$1; shift
mkdir -p gpurun_out
python -m pytest tests/test_gpu_api.py -m gpu -q -k pool 2>&1 | tail -2 > gpurun_out/${T}_pool_pytest.log
last=1
for N in "$@"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 295$N bench.py --gpus $N --steps 20 --warmup 3 --no-compressed > gpurun_out/${T}_bench_${N}gpu.json 2> gpurun_out/${T}_bench_${N}gpu.err
  last=$N
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node $last --master-addr 127.0.0.1 --master-port 29599 bench.py --gpus $last --workload cfg5 --steps 20 --warmup 3 --no-compressed > gpurun_out/${T}_bench_cfg5_${last}gpu.json 2> gpurun_out/${T}_bench_cfg5_${last}gpu.err
python bench.py --impl reference --workload cfg5 --steps 6 --warmup 1 > gpurun_out/${T}_bench_cfg5_reference.json 2>/dev/null
cat gpurun_out/${T}_pool_pytest.log
python - "$T" "$@" <<'PY'
import json, sys
T = sys.argv[1]
for n in sys.argv[2:] + ["cfg5_" + sys.argv[-1]]:
    try:
        d = json.loads(open(f"gpurun_out/{T}_bench_{n}gpu.json").read().strip().splitlines()[-1])
        e = d["e2e"]
        print(n, "gpus: value", round(d["value"]), "e2e", round(e["value"], 1), "ceiling", round(e.get("host_ceiling_fps", 0), 1), "of ceiling", round(e.get("of_host_ceiling", 0), 3), "cpus/rank", d["run"]["host_cpus_per_rank"])
    except Exception as ex:
        print(n, "failed", ex)
PY
