L=$PWD/raysnail_b200/libraysnail_b200_qf.so
RAYSNAIL_B200_LIB=$L timeout 600 python -m pytest tests -m gpu -q > gpurun_out/t_qf.log 2>&1; tail -3 gpurun_out/t_qf.log
run() { timeout 200 python bench.py --no-cpu "$@" 2>/dev/null | tail -1; }
for v in base qf; do
  if [ $v = qf ]; then export RAYSNAIL_B200_LIB=$L; fi
  run --scene quadric --width 1920 --height 1080 --sqrt-spp 16 --steps 3 --warmup 3 > gpurun_out/ab_${v}_quadric.json
  run --scene cornell --width 3840 --height 2160 --sqrt-spp 16 --steps 2 --warmup 3 > gpurun_out/ab_${v}_cornell.json
  run --scene example --width 800 --height 500 --sqrt-spp 8 --steps 5 --warmup 3 > gpurun_out/ab_${v}_example.json
  run --scene declares --width 1920 --height 1080 --sqrt-spp 8 --steps 3 --warmup 3 > gpurun_out/ab_${v}_declares.json
done
