#!/bin/bash
# usage (on the GPU box): tools/prof_lens.sh <tag>  -> gpurun_out/prof_<tag>_lens.ncu-rep
# One `ncu --set full` capture of the lens launch (stitch_fast2_kernel<true>: cv::undistort of all cameras' frames)
# in the middle of a short device-resident `bench.py --lens` run, only after the same command exited 0 without ncu.
T=$1; shift
CMD="python bench.py --lens --steps 2 --warmup 3 --frames-per-step 16 --no-e2e --no-cpu-baseline --no-host-ceiling --no-compressed $*"
BS_DEBUG_PLAN_SUMMARY=1 $CMD > gpurun_out/plain_${T}_lens.log 2>&1 && ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:stitch_fast2_kernelILb1 -s 20 -c 1 -o gpurun_out/prof_${T}_lens -f $CMD > gpurun_out/ncu_${T}_lens.log 2>&1
grep "lens plan" gpurun_out/plain_${T}_lens.log | head -4; tail -1 gpurun_out/ncu_${T}_lens.log
