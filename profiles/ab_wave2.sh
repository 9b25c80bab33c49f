for w in 67108864 134217728; do for st in 2 4; do
python bench.py --steps 3 --warmup 2 --sqrt-spp 16 --no-cpu --no-others --wave $w --wave-streams $st 2>&1 | tail -1 | python -c "
import json,sys
try:
  d=json.loads(sys.stdin.read()); print('wave $w streams $st', round(d['value']/1e6,1), 'Msamples/s', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']/1e6,1))
except Exception as e: print('wave $w streams $st FAILED', e)"
done; done
python bench.py --steps 2 --warmup 2 --no-cpu --no-others --scene mesh_blinn:708 --width 2048 --height 2048 --sqrt-spp 8 --wave 134217728 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('mesh wave 128Mi', round(d['value']/1e6,1), 'Msamples/s', round(d['ms_per_step'],2))"
python bench.py --steps 2 --warmup 2 --no-cpu --no-others --scene cornell --width 3840 --height 2160 --sqrt-spp 16 --wave 67108864 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('cornell wave 64Mi', round(d['value']/1e6,1), 'Msamples/s', round(d['ms_per_step'],2))"
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
