#!/bin/bash
# Round evidence on one B200 (run under gpurun): bench lines, ncu launch list, ncu --set full of one whole
# timestep, resident-kernel captures. Everything lands in gpurun_out/; tools/ncu_summary.py turns it into profiles/.
set -u
R=${1:-r2}
export FLUIDB200_CPU_GRID=${FLUIDB200_CPU_GRID:-1536}
timeout 600 python bench.py > gpurun_out/bench_$R.json 2> gpurun_out/bench_$R.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_arm_$R.json 2>/dev/null; echo "ref rc=$?"
# the profiled command: the headline step only (the side measurements are switched off by their knobs)
export FLUIDB200_SKIP_SMALL=1 FLUIDB200_SKIP_GENERAL=1 FLUIDB200_SKIP_E2E_PAIR=1 FLUIDB200_E2E_STEPS=1 FLUIDB200_CPU_GRID=64
CMD="python bench.py --steps 2 --warmup 3"
timeout 300 $CMD > gpurun_out/plain_$R.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$R.csv $CMD > gpurun_out/ncu_launches_$R.log 2>&1
echo "launch list rc=$?"
# (a timestep is 41 kernels: predictor, 13 x {edge, wall, fast} -- 9 launches of depth 8, then 4 of depth 7 --, corrector;
#  124 kernels run before the first timed step: init + 3 warm-up steps)
timeout 300 $CMD > gpurun_out/plain2_$R.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:'jacobi|predictor|corrector' -s 123 -c 4 -o gpurun_out/prof_step_a_$R $CMD > gpurun_out/ncu_full_a_$R.log 2>&1
echo "full capture a rc=$?"
timeout 300 $CMD > gpurun_out/plain3_$R.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:'jacobi|predictor|corrector' -s 160 -c 4 -o gpurun_out/prof_step_b_$R $CMD > gpurun_out/ncu_full_b_$R.log 2>&1
echo "full capture b rc=$?"
# (the reports themselves are ~9 MB per kernel: only their text pages travel back)
for x in a b; do
  ncu -i gpurun_out/prof_step_${x}_$R.ncu-rep --page raw --csv > gpurun_out/prof_step_${x}_$R.raw.csv 2>/dev/null
  ncu -i gpurun_out/prof_step_${x}_$R.ncu-rep --page details > gpurun_out/prof_step_${x}_$R.details.txt 2>/dev/null
  rm -f gpurun_out/prof_step_${x}_$R.ncu-rep
done
for n in 41 1000; do
  timeout 120 python tools/run_once.py $n 20 50 > gpurun_out/plain_res$n.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:resident -c 1 -o gpurun_out/prof_resident${n}_$R python tools/run_once.py $n 20 50 > gpurun_out/ncu_res$n.log 2>&1
  echo "resident $n rc=$?"
  ncu -i gpurun_out/prof_resident${n}_$R.ncu-rep --page raw --csv > gpurun_out/prof_resident${n}_$R.raw.csv 2>/dev/null
  ncu -i gpurun_out/prof_resident${n}_$R.ncu-rep --page source --csv > gpurun_out/prof_resident${n}_$R.source.csv 2>/dev/null
  ncu -i gpurun_out/prof_resident${n}_$R.ncu-rep --page details > gpurun_out/prof_resident${n}_$R.details.txt 2>/dev/null
  rm -f gpurun_out/prof_resident${n}_$R.ncu-rep
done
du -sh gpurun_out
