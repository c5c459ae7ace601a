#!/bin/bash
# predictor / corrector ms at 16384^2 by rows per chunk (FLUIDB200_RPC applies to every kernel of the step; only these two columns matter here)
for r in "$@"; do
  o=$(FLUIDB200_RPC=$r timeout 100 python tools/phase_times.py 16384 16 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['predictor_ms'], d['corrector_ms'])")
  echo "rpc=$r $o"
done
