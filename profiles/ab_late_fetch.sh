#!/bin/bash
# A/B of RS_TUNE_FIRST_LATE_FETCH (make VARIANT=_lf TUNE="-DRS_TUNE_FIRST_LATE_FETCH=1") on one B200
L=$PWD/raysnail_b200/libraysnail_b200_lf.so
RAYSNAIL_B200_LIB=$L timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t_lf.log 2>&1; tail -2 gpurun_out/t_lf.log
run() { timeout 200 python bench.py --no-cpu "$@" 2>/dev/null | tail -1; }
for rep in 1 2; do
for v in base lf; do
  if [ $v = lf ]; then export RAYSNAIL_B200_LIB=$L; else unset RAYSNAIL_B200_LIB; fi
  run --sqrt-spp 16 --steps 3 --warmup 3 > gpurun_out/ablf_${v}_rtow_$rep.json
  run --scene cornell --width 3840 --height 2160 --sqrt-spp 8 --steps 3 --warmup 3 > gpurun_out/ablf_${v}_cornell_$rep.json
  run --scene quadric --width 1920 --height 1080 --sqrt-spp 8 --steps 3 --warmup 3 > gpurun_out/ablf_${v}_quadric_$rep.json
done
done
