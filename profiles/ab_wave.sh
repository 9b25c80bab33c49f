for w in 16777216 33554432 67108864; do for st in 2 4; do
python bench.py --steps 3 --warmup 2 --sqrt-spp 16 --no-cpu --no-others --wave $w --wave-streams $st 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('wave $w streams $st', round(d['value']/1e6,1), 'Msamples/s', round(d['ms_per_step'],2))"
done; done
for w in 33554432 67108864; do
python bench.py --steps 2 --warmup 2 --no-cpu --no-others --scene mesh_blinn:708 --width 2048 --height 2048 --sqrt-spp 8 --wave $w 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('mesh wave $w', round(d['value']/1e6,1), 'Msamples/s', round(d['ms_per_step'],2))"
done
