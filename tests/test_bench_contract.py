"""The benchmark's JSON contract, as far as it can be checked without a GPU: the reference arm
(`bench.py --impl reference`) runs the OpenCV CPU path and prints one JSON line with the keys the
driver reads; the B200 arm refuses to run without a device instead of falling back to the CPU."""
import json
import os
import subprocess
import sys

from conftest import ROOT

REQUIRED = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
            "vs_baseline", "dtype", "data", "config", "e2e", "cpu_baseline", "impl"}


def run(*args, timeout=600):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          timeout=timeout, env=env, cwd=ROOT)


def test_reference_arm_prints_the_contract_line():
    out = run("--impl", "reference", "--workload", "cfg2", "--steps", "2", "--warmup", "1")
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert REQUIRED <= set(line), REQUIRED - set(line)
    assert line["impl"] == "reference" and line["metric"] == "panorama_frames_per_sec" and line["unit"] == "frames/s"
    assert line["value"] > 0 and line["higher_is_better"] is True and line["dtype"] == "u8"
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in line["config"]


def test_b200_arm_has_no_cpu_fallback():
    out = run("--workload", "cfg2", "--steps", "1", "--warmup", "1", "--no-cpu-baseline", timeout=300)
    assert out.returncode != 0
    assert "no CUDA device" in (out.stderr + out.stdout)
