#!/bin/bash
# us per depth-8 (and 7) Jacobi launch by the number of row chunks (no wave alignment): tools/scan_chunks.sh NX NY c1 c2 ...
NX=$1; NY=$2; shift 2
W=$(( (NY + 119) / 120 ))   # strips of 120 owned columns at depth 8 (approximate: only scales the task target)
for c in "$@"; do
  t=$((c*W))
  r=$(NX=$NX NY=$NY FLUIDB200_NO_WAVE_ALIGN=1 FLUIDB200_TARGET_TASKS=$t timeout 100 python tools/jacobi_depth.py 7 8 2>&1 | tail -1)
  echo "chunks~$c $r"
done
