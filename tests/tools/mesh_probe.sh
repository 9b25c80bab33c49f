#!/bin/bash
# mesh path evidence: bench + per-ray cost + ncu metrics of extend<1>
python bench.py --steps 2 --warmup 2 --no-cpu --no-others --scene mesh_blinn:708 --width 2048 --height 2048 --sqrt-spp 8 2>&1 | tail -1 > gpurun_out/mesh_bench.json
cat gpurun_out/mesh_bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']/1e6,1),'Msamples/s', d['ms_per_step'], d['roofline']['kernel_ms_per_step'])"
RAYSNAIL_B200_LIB=$PWD/raysnail_b200/libraysnail_b200_cost.so python profiles/cost_dump.py mesh_blinn:708 2048 2048 > gpurun_out/mesh_cost.txt 2>&1
grep -v "sorted by" gpurun_out/mesh_cost.txt | head -40
ncu --clock-control none -k regex:"extend_kernel" --launch-skip 8 --launch-count 3 --section SpeedOfLight --section MemoryWorkloadAnalysis --section WarpStateStats --section SchedulerStats --section Occupancy --section LaunchStats \
  python bench.py --steps 1 --warmup 1 --no-cpu --no-others --scene mesh_blinn:708 --width 2048 --height 2048 --sqrt-spp 2 > gpurun_out/mesh_ncu.txt 2>&1
grep -E "extend_kernel|Duration|DRAM Throughput|L1/TEX Hit|L2 Hit|Registers|Achieved Occ|Issue Slots Busy|No Eligible|Mem Busy|Max Bandwidth|Local|Theoretical Occ|Compute \(SM\)|Memory Throughput" gpurun_out/mesh_ncu.txt | head -60
