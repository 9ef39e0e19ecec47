#!/bin/bash
# usage (on the GPU box): tools/prof.sh <tag> [extra bench.py args]  -> gpurun_out/prof_<tag>.ncu-rep
# One `ncu --set full` capture of the frame kernel in the middle of a short device-resident bench run,
# only after the same command has exited 0 without ncu.
T=$1; shift
CMD="python bench.py --steps 2 --warmup 3 --frames-per-step 16 --no-e2e --no-cpu-baseline $*"
$CMD > gpurun_out/plain_$T.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:stitch_fast -s 30 -c 1 -o gpurun_out/prof_$T -f $CMD > gpurun_out/ncu_$T.log 2>&1
tail -1 gpurun_out/plain_$T.log; tail -1 gpurun_out/ncu_$T.log
