#!/bin/bash
# quick A/B of library variants on the headline scene at 64 spp: bash profiles/quick.sh "" _t1 _t2
for v in "$@"; do
  L=raysnail_b200/libraysnail_b200$v.so
  RAYSNAIL_B200_LIB=$PWD/$L python bench.py --steps 3 --warmup 2 --sqrt-spp 8 --no-cpu ${QUICK_ARGS} 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms_per_step']
print('variant [$v]', round(d['value']/1e6,1), 'Msamples/s', round(d['mrays_per_s'],1), 'Mrays/s', {a:round(b,2) for a,b in k.items()})"
done
