#!/bin/bash
# bench.py at N ranks on the GPUs of this box (the line carries the weak series as value and the strong one in config.strong)
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3 2> gpurun_out/scale_$N.err | tail -1 > gpurun_out/scale_$N.json
python - <<PY
import json
d=json.load(open("gpurun_out/scale_$N.json"))
print($N, "weak", round(d["value"],2), "Gcell/s", round(d["ms_per_step"],3), "ms; strong", json.dumps(d["config"]["strong"]), "e2e ms", round(d["e2e"]["ms_per_step"],1), d["e2e"].get("host_link_GBps_aggregate"))
PY
