#!/bin/bash
# us per Jacobi launch by rows per chunk: tools/scan_rpc.sh NX NY depth rpc1 rpc2 ...
NX=$1; NY=$2; K=$3; shift 3
for r in "$@"; do
  o=$(NX=$NX NY=$NY FLUIDB200_RPC=$r timeout 100 python tools/jacobi_depth.py $K 2>&1 | tail -1)
  echo "rpc=$r chunks=$(( (NX - 27 + r - 1) / r )) $o"
done
