OUT=gpurun_out/r02h_scale48.jsonl
: > $OUT
for N in 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+N)) bench.py --gpus $N --steps 3 --warmup 3 --no-cpu --no-others 2>>gpurun_out/r02h_scale48.err | tail -1 >> $OUT
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu --no-others --scene example --width 800 --height 500 --sqrt-spp 8 2>>gpurun_out/r02h_scale48.err | tail -1 >> $OUT
python - "$OUT" <<'PY'
import json, sys
for line in open(sys.argv[1]):
    try: d = json.loads(line)
    except Exception: continue
    print(d["n_gpus"], d["config"]["workload"][:30], round(d["value"] / 1e9, 3), "G samples/s", round(d["ms_per_step"], 2), "ms/step e2e", round(d["e2e"]["value"] / 1e9, 3))
PY
