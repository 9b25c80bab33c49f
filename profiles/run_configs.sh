#!/bin/bash
# Runs every BASELINE.json config on one B200 and stores one JSON line per config.
# usage (on the GPU box): bash profiles/run_configs.sh gpurun_out/configs.jsonl
OUT=${1:-gpurun_out/configs.jsonl}
: > $OUT
run() { python bench.py --no-cpu "$@" 2>>gpurun_out/configs.err | tail -1 >> $OUT; }
run --scene example    --width 800  --height 500  --sqrt-spp 8  --steps 5 --warmup 3
run --scene rtow_final --width 1920 --height 1080 --sqrt-spp 32 --steps 3 --warmup 3
run --scene quadric    --width 1920 --height 1080 --sqrt-spp 16 --steps 3 --warmup 3
run --scene csg        --width 1920 --height 1080 --sqrt-spp 16 --steps 3 --warmup 3
run --scene mesh_blinn:708 --width 2048 --height 2048 --sqrt-spp 8 --steps 3 --warmup 3
run --scene cornell    --width 3840 --height 2160 --sqrt-spp 16 --steps 2 --warmup 3
python - "$OUT" <<'PY'
import json, sys
for line in open(sys.argv[1]):
    try: d = json.loads(line)
    except Exception: continue
    r = d["roofline"]
    print(f'{d["config"]["workload"][:58]:58s} {d["value"]/1e6:9.1f} Msamples/s {d["mrays_per_s"]:8.1f} Mrays/s  {d["ms_per_step"]:9.1f} ms/step  seg/sample {d["segments_per_sample"]:.2f}  dom {r["kernel"]} frac {r["frac"]:.4f} pipeline frac {r["pipeline"]["frac"]:.4f}')
PY
