#!/bin/bash
# 1 -> 8 GPU scaling of the headline workload, C1 (800x500x64 spp: the launch-bound end) and BASELINE config 5 (Cornell box,
# 3840x2160, 4096 spp) on 8 GPUs, the multi-device GPU tests and the torchrun image check.
# usage (on an 8-GPU box): bash profiles/scale.sh r02c
TAG=${1:-rXX}
OUT=gpurun_out/${TAG}_scale.jsonl
ERR=gpurun_out/${TAG}_scale.err
: > $OUT
(timeout 600 python -m pytest tests -m gpu -q -k "rs_multi or fused_peer" 2>&1 | tail -2) | tee gpurun_out/${TAG}_n8_tests.txt
python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu --no-others 2>>$ERR | tail -1 >> $OUT
P=29600
for N in 2 4 8; do
  P=$((P+1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N --steps 3 --warmup 3 --no-cpu --no-others 2>>$ERR | tail -1 >> $OUT
done
python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu --no-others --scene example --width 800 --height 500 --sqrt-spp 8 2>>$ERR | tail -1 >> $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29609 bench.py --gpus 8 --steps 5 --warmup 3 --no-cpu --no-others --scene example --width 800 --height 500 --sqrt-spp 8 2>>$ERR | tail -1 >> $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29610 bench.py --gpus 8 --steps 1 --warmup 3 --no-cpu --no-others --scene cornell --width 3840 --height 2160 --sqrt-spp 64 2>>$ERR | tail -1 >> $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 tests/multi_gpu_check.py 2>>$ERR | tail -1 >> $OUT
python - "$OUT" <<'PY'
import json, sys
for line in open(sys.argv[1]):
    try: d = json.loads(line)
    except Exception: continue
    if "value" in d: print(d["n_gpus"], d["config"]["workload"][:40], round(d["value"] / 1e9, 3), "G samples/s", round(d["ms_per_step"], 2), "ms/step e2e", round(d["e2e"]["value"] / 1e9, 3))
    else: print(d)
PY
