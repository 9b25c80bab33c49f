#!/bin/bash
# 2-GPU evidence (on a 2-GPU box): the GPU tests that need two devices (rs_multi_render on devices (0, 1), fused peer
# finalize), the torchrun image check, and the headline + C1 bench lines at N = 1 and 2.
# usage: bash profiles/scale2.sh r02c
TAG=${1:-rXX}
(timeout 600 python -m pytest tests -m gpu -q -k "rs_multi or fused_peer or ipc" 2>&1 | tail -3) | tee gpurun_out/${TAG}_n2_tests.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 tests/multi_gpu_check.py 2>>gpurun_out/${TAG}_n2.err | tail -1 | tee gpurun_out/${TAG}_n2_check.json
OUT=gpurun_out/${TAG}_scale2.jsonl
: > $OUT
python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu --no-others 2>>gpurun_out/${TAG}_n2.err | tail -1 >> $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29622 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu --no-others 2>>gpurun_out/${TAG}_n2.err | tail -1 >> $OUT
python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu --no-others --scene example --width 800 --height 500 --sqrt-spp 8 2>>gpurun_out/${TAG}_n2.err | tail -1 >> $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29623 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu --no-others --scene example --width 800 --height 500 --sqrt-spp 8 2>>gpurun_out/${TAG}_n2.err | tail -1 >> $OUT
python - "$OUT" <<'PY'
import json, sys
for line in open(sys.argv[1]):
    try: d = json.loads(line)
    except Exception: continue
    if "value" in d: print(d["n_gpus"], d["config"]["workload"][:40], round(d["value"] / 1e9, 3), "G samples/s", round(d["ms_per_step"], 2), "ms/step e2e", round(d["e2e"]["value"] / 1e9, 3))
    else: print(d)
PY
