#!/bin/bash
# ms per end-to-end step (fluidb200_migrate_solve, 16384^2) with and without the tapered last chunk, interleaved
export FLUIDB200_SKIP_SMALL=1 FLUIDB200_SKIP_GENERAL=1 FLUIDB200_SKIP_E2E_PAIR=1 FLUIDB200_CPU_GRID=64 FLUIDB200_E2E_STEPS=5
for rep in 1 2; do for t in 0 1; do
  FLUIDB200_E2E_TAPER=$t timeout 200 python bench.py --steps 2 --warmup 3 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('taper', $t, round(d['e2e']['ms_per_step'],2), round(d['e2e']['link_GBps_each_way'],1))"
done; done
