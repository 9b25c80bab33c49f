#!/bin/bash
# ncu --set full of the non-headline tiers (on the GPU box): bounce 1 extend of the 1 M-triangle mesh (C4) and
# two bounces of sdl/quadric.sdl (C3a, generic tier).  usage: bash profiles/capture_others.sh r02x
TAG=${1:-rXX}
ncu --set full --import-source on --clock-control none -k regex:"extend_kernel|extend_stream_kernel" --launch-skip 8 --launch-count 3 \
  -o gpurun_out/${TAG}_mesh -f python bench.py --steps 1 --warmup 1 --no-cpu --no-others --scene mesh_blinn:708 --width 2048 --height 2048 --sqrt-spp 2 > gpurun_out/${TAG}_mesh.log 2>&1
python profiles/ncu_metrics.py gpurun_out/${TAG}_mesh.ncu-rep > gpurun_out/${TAG}_mesh_ncu_summary.txt 2>&1
ncu --set full --import-source on --clock-control none -k regex:"shade_kernel|extend_kernel" --launch-skip 24 --launch-count 6 \
  -o gpurun_out/${TAG}_quadric -f python bench.py --steps 1 --warmup 1 --no-cpu --no-others --scene quadric --sqrt-spp 2 > gpurun_out/${TAG}_quadric.log 2>&1
python profiles/ncu_metrics.py gpurun_out/${TAG}_quadric.ncu-rep > gpurun_out/${TAG}_quadric_ncu_summary.txt 2>&1
