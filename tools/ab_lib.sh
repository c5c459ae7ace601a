#!/bin/bash
# A/B of two builds on one box: tools/ab_lib.sh path/to/variant.so   (jacobi launch times for the tall and the slab shape, whole step)
V=$1
for lib in default $V; do
  if [ $lib = default ]; then unset FLUIDB200_LIB; else export FLUIDB200_LIB=$PWD/$lib; fi
  echo "== $lib"
  timeout 100 python tools/jacobi_depth.py 7 8 2>&1 | tail -1
  NX=4096 NY=32768 timeout 100 python tools/jacobi_depth.py 7 8 2>&1 | tail -1
  timeout 100 python tools/phase_times.py 16384 100 2>&1 | tail -1
done
