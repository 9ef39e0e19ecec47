#!/bin/bash
# SASS opcode histogram of the per-frame kernels in the shipped library -> profiles/<tag>_sass_opcodes.txt
# (shows the Blackwell-native instructions: UTMALDG = TMA tensor load, UBLKCP = bulk copy, SYNCS = mbarrier,
#  IDP.2A = dp2a, FFMA2/FADD2 = packed fp32)
T=${1:-r02}
OUT=profiles/${T}_sass_opcodes.txt
: > $OUT
for k in stitch_fast2_kernel stitch_fast_kernel undistort_remap_kernel plan_kernel; do
  echo "== $k" >> $OUT
  cuobjdump -sass image-stitching_b200/libb200stitch.so | awk -v k=$k '/Function : /{p=index($0,k)>0} p' | grep -E "^\s+/\*[0-9a-f]{4,5}\*/" | sed -E 's#^\s+/\*[0-9a-f]+\*/\s+(@!?U?P[0-9T]+ +)?##' | awk '{print $1}' | sed -E 's/\..*//' | sort | uniq -c | sort -rn | awk '{printf "%s:%s ", $2, $1} END{print ""}' >> $OUT
  cuobjdump -sass image-stitching_b200/libb200stitch.so | awk -v k=$k '/Function : /{p=index($0,k)>0} p' | grep -oE "UTMALDG[.A-Z0-9]*|UBLKCP[.A-Z0-9]*|SYNCS[.A-Z0-9]*|IDP[.A-Z0-9]*|FFMA2|FADD2|UTMASTG[.A-Z0-9]*" | sort | uniq -c | awk '{printf "   %s x%s", $2, $1} END{print ""}' >> $OUT
done
cat $OUT
