#!/bin/bash
# 1 -> 8 GPU scaling of the headline workload + BASELINE config 5 (Cornell box, 3840x2160, 4096 spp) on 8 GPUs.
# usage (on an 8-GPU box): bash profiles/scale.sh gpurun_out/r01g_scale.jsonl
OUT=${1:-gpurun_out/scale.jsonl}
: > $OUT
python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu 2>>gpurun_out/scale.err | tail -1 >> $OUT
P=29600
for N in 2 4 8; do
  P=$((P+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 3 --warmup 3 --no-cpu 2>>gpurun_out/scale.err | tail -1 >> $OUT
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29610 bench.py --gpus 8 --steps 1 --warmup 3 --no-cpu --scene cornell --width 3840 --height 2160 --sqrt-spp 64 2>>gpurun_out/scale.err | tail -1 >> $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 tests/multi_gpu_check.py 2>>gpurun_out/scale.err | tail -1 >> $OUT
python - "$OUT" <<'PY'
import json, sys
for line in open(sys.argv[1]):
    try: d = json.loads(line)
    except Exception: continue
    if "value" in d: print(d["n_gpus"], d["config"]["workload"][:40], round(d["value"] / 1e9, 3), "G samples/s", round(d["ms_per_step"], 1), "ms/step e2e", round(d["e2e"]["value"] / 1e9, 3))
    else: print(d)
PY
