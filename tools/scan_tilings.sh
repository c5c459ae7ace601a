#!/bin/bash
# ms per 1000 steps x 50 sweeps of the resident solver by forced tiling (R = 8): tools/scan_tilings.sh NX NY "wx wy" ...
NX=$1; NY=$2; shift 2
echo "default $(NY=$NY S=1000 timeout 100 python tools/bench_resident.py $NX 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['resident_ms'], d['took_resident_path'])")"
for t in "$@"; do set -- $t
  echo "wx=$1 wy=$2 $(NY=$NY FLUIDB200_RESIDENT_R=8 FLUIDB200_RESIDENT_WX=$1 FLUIDB200_RESIDENT_WY=$2 S=1000 timeout 100 python tools/bench_resident.py $NX 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['resident_ms'], d['took_resident_path'])")"
done
