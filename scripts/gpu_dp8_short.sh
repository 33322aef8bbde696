#!/bin/bash
# 8 GPUs, short: AlexNet at N=8, then N=4 on each half of the box side by side (second half: the NCCL arm)
mkdir -p gpurun_out
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
show() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], d.get("impl"), d["n_gpus"], round(d["value"]), round(d["ms_per_step"], 3), (d.get("e2e") or {}).get("value"))
except Exception as e:
    print(sys.argv[1], "parse failed", e)
PY
}
timeout 200 $TR8 --master-port 29841 bench.py --gpus 8 --steps 30 --warmup 5 > gpurun_out/a8_final.json 2> gpurun_out/a8_final.err; echo a8 rc=$?; show gpurun_out/a8_final.json
(CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 200 $TR4 --master-port 29842 bench.py --gpus 4 --steps 30 --warmup 5 --skip-e2e > gpurun_out/a4_final.json 2> gpurun_out/a4_final.err) &
(CUDA_VISIBLE_DEVICES=4,5,6,7 timeout 200 $TR4 --master-port 29843 bench.py --gpus 4 --steps 30 --warmup 5 --skip-e2e --impl nccl > gpurun_out/a4_nccl.json 2> gpurun_out/a4_nccl.err) &
wait; show gpurun_out/a4_final.json; show gpurun_out/a4_nccl.json
