for c in 8 9 10 11 12 13 14 15 16 17 18 19 20 22 24 26 28 32 36 40; do
  t=$((c*274))
  r=$(NX=4096 NY=32768 FLUIDB200_NO_WAVE_ALIGN=1 FLUIDB200_TARGET_TASKS=$t timeout 100 python tools/jacobi_depth.py 8 2>&1 | tail -1)
  echo "chunks~$c $r"
done
