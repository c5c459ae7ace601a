#!/bin/bash
# ms per end-to-end step (fluidb200_migrate_solve, 16384^2) by rows per pipeline chunk
export FLUIDB200_SKIP_SMALL=1 FLUIDB200_SKIP_GENERAL=1 FLUIDB200_SKIP_E2E_PAIR=1 FLUIDB200_CPU_GRID=64 FLUIDB200_E2E_STEPS=5
for c in ${@:-384 512 640 768 1024}; do
  FLUIDB200_E2E_CHUNK=$c timeout 200 python bench.py --steps 2 --warmup 3 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print($c, round(d['e2e']['ms_per_step'],2), round(d['e2e']['link_GBps_each_way'],1))"
done
