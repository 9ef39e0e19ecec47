#!/bin/bash
# usage (on the GPU box): tools/jpeg_launches.sh <tag>  -> gpurun_out/<tag>_jpeg_launches.csv
# Durations of the JPEG decoder's kernels (ncu launch list) while bench.py runs its compressed records, after the same
# command has exited 0 without ncu.
T=$1
CMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-host-ceiling"
$CMD > gpurun_out/${T}_jpeg_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:jpeg_ -s 400 -c 600 --csv --log-file gpurun_out/${T}_jpeg_launches.csv $CMD > gpurun_out/${T}_jpeg_ncu.log 2>&1
tail -1 gpurun_out/${T}_jpeg_ncu.log | cut -c1-200
