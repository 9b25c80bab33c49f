#!/bin/bash
# mesh A/B: bash tests/tools/mesh_ab.sh "label|ENV=..|_variant" ...
for spec in "$@"; do
  IFS='|' read -r label envs variant <<< "$spec"
  L=$PWD/raysnail_b200/libraysnail_b200$variant.so
  env $envs RAYSNAIL_B200_LIB=$L python bench.py --steps 2 --warmup 2 --no-cpu --no-others --scene mesh_blinn:708 --width 2048 --height 2048 --sqrt-spp 8 2>&1 | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms_per_step']
    print('[mesh $label]', round(d['value']/1e6,1), 'Msamples/s', round(d['ms_per_step'],2), 'ms/step', {a:round(b,2) for a,b in k.items()})
except Exception as e: print('[mesh $label] FAILED', e)"
done
