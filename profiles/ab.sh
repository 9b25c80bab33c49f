#!/bin/bash
# A/B of (env, library variant) pairs on the headline scene at 64 spp.
# usage: bash profiles/ab.sh "label|ENV=1 ENV2=2|_variant|extra bench args" ...
for spec in "$@"; do
  IFS='|' read -r label envs variant extra <<< "$spec"
  L=$PWD/raysnail_b200/libraysnail_b200$variant.so
  env $envs RAYSNAIL_B200_LIB=$L python bench.py --steps 3 --warmup 2 --sqrt-spp 8 --no-cpu --no-others $extra 2>&1 | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms_per_step']
    print('[$label]', round(d['value']/1e6,1), 'Msamples/s', round(d['ms_per_step'],2), 'ms/step', {a:round(b,2) for a,b in k.items()})
except Exception as e: print('[$label] FAILED', e)"
done
