#!/bin/bash
# 4 GPUs, 4096 rows per GPU (the per-GPU shape of the 8-GPU strong-scaling run): effect of the slab driver's stream choices
export FLUIDB200_E2E_STEPS=1 FLUIDB200_CPU_GRID=128 FLUIDB200_SKIP_SMALL=1 FLUIDB200_SCALING=strong FLUIDB200_STRONG_ROWS=16384
N=${N:-4}
run() { echo "== $*"; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 5 --warmup 3 2>/dev/null | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['config']['workload'][:60], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value'],1))"; }
run A=1
run FLUIDB200_ONE_EDGE_STREAM=1
run FLUIDB200_NCCL_LOWPRIO=1
run FLUIDB200_ONE_EDGE_STREAM=1 FLUIDB200_NCCL_LOWPRIO=1
