#!/bin/bash
# Weak- and strong-scaling series on one 8-GPU box (results appended to gpurun_out/scaling.jsonl).
#   weak   : 16384 rows per GPU (BASELINE.json configs[4]), N in $WEAK_N
#   strong : 32768 x 32768 total (configs[3]), N in $STRONG_N
set -u
OUT=gpurun_out/scaling.jsonl; mkdir -p gpurun_out; : > $OUT
export FLUIDB200_E2E_STEPS=1 FLUIDB200_CPU_GRID=128 FLUIDB200_SKIP_SMALL=1
STEPS=${STEPS:-5}
run() { # N, extra env...
  local n=$1; shift
  if [ "$n" = 1 ]; then env "$@" python bench.py --gpus 1 --steps $STEPS --warmup 3 2>gpurun_out/scaling_err_$n.log | tail -1 >> $OUT
  else env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --steps $STEPS --warmup 3 2>gpurun_out/scaling_err_$n.log | tail -1 >> $OUT; fi
}
for n in ${WEAK_N:-1 2 4 8}; do run $n FLUIDB200_SCALING=weak; done
for n in ${STRONG_N:-1 2 4 8}; do
  if [ "$n" = 1 ]; then run 1 FLUIDB200_BENCH_GRID=32768; else run $n FLUIDB200_SCALING=strong; fi
done
python - <<'PY'
import json
for l in open("gpurun_out/scaling.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line:", l[:200]); continue
    print(d["n_gpus"], d["scaling"], d["config"]["workload"][:60], "value", round(d["value"], 2), "ms/step", round(d["ms_per_step"], 2),
          "e2e", round(d["e2e"]["value"], 2))
PY
