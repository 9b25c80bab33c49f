#!/bin/bash
# per-scene A/B at the default frame size: bash tests/tools/ab_scenes.sh "label|ENV=1 ENV2=2" ...
SCENES="${SCENES:-quadric csg cornell declares example all_feature}"
for spec in "$@"; do
  IFS='|' read -r label envs variant <<< "$spec"
  L=$PWD/raysnail_b200/libraysnail_b200$variant.so
  for sc in $SCENES; do
    env $envs RAYSNAIL_B200_LIB=$L python bench.py --steps 3 --warmup 2 --sqrt-spp 8 --no-cpu --no-others --scene $sc 2>&1 | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms_per_step']
    print('[$sc $label]', round(d['value']/1e6,1), 'Msamples/s', round(d['ms_per_step'],2), 'ms/step', {a:round(b,2) for a,b in k.items()})
except Exception as e: print('[$sc $label] FAILED', e)"
  done
done
