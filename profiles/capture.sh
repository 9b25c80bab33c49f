#!/bin/bash
# Round evidence on one B200: the bench line, the ncu launch list of the same command and one
# `ncu --set full` capture of the two hot kernels (bounce 0 and bounce 1 of the second wave).
# usage (on the GPU box): bash profiles/capture.sh r02x <commit>   (the commit is recorded in the traffic json)
TAG=${1:-rXX}
COMMIT=${2:-unknown}
set -x
python bench.py --steps 3 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv \
    python bench.py --steps 1 --warmup 1 --sqrt-spp 4 --no-cpu > gpurun_out/${TAG}_launches.log 2>&1
# at 16 spp the frame is ONE wave of 1920x1080x16 = 33 177 600 paths (64 Mi paths per wave is the default)
# one wave = 8 bounces x (extend, shade<Lambertian>, shade<rest>) = 24 matching launches (bounce 0 = the extend
# that generates the camera rays itself): skip the first wave, capture bounce 0 and bounce 1 of the second one
ncu --set full --import-source on --clock-control none -k regex:"shade_kernel|extend_kernel" --launch-skip 24 --launch-count 6 \
    -o gpurun_out/${TAG}_full python bench.py --steps 1 --warmup 1 --sqrt-spp 4 --no-cpu > gpurun_out/${TAG}_full.log 2>&1
python profiles/ncu_metrics.py gpurun_out/${TAG}_full.ncu-rep --traffic-json gpurun_out/${TAG}_traffic.json --paths 33177600 --commit $COMMIT > gpurun_out/${TAG}_ncu_summary.txt 2>&1
tail -1 gpurun_out/${TAG}_bench.json | cut -c1-400
