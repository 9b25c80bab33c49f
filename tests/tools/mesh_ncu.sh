#!/bin/bash
# full ncu capture of the bounce-1 extend kernel on the 1 M-triangle mesh: batch and streaming forms
for mode in 0 1; do
  RAYSNAIL_B200_STREAM=$mode ncu --set full --import-source on --clock-control none -k regex:"extend_kernel|extend_stream_kernel" --launch-skip 9 --launch-count 1 \
    -o gpurun_out/r02_mesh_b1_stream$mode -f python bench.py --steps 1 --warmup 1 --no-cpu --no-others --scene mesh_blinn:708 --width 2048 --height 2048 --sqrt-spp 2 > gpurun_out/mesh_ncu_$mode.log 2>&1
  tail -2 gpurun_out/mesh_ncu_$mode.log | cut -c1-300
done
