#!/bin/bash
# 2 GPUs: AlexNet trajectory check (fused + nccl exchange) and the bench pair
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29731 scripts/dp_check.py alexnet fused,nccl > gpurun_out/dp_check_alexnet2.log 2>&1; echo dpcheck-alexnet rc=$?; grep -E "DP_CHECK|differ|Error|error" gpurun_out/dp_check_alexnet2.log | cut -c1-200 | head -12
timeout 200 $TR --master-port 29732 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/b2.json 2> gpurun_out/b2.err; echo bench2 rc=$?
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/b2.json").read().strip().splitlines()[-1]); print(d["value"], d["ms_per_step"], d["launches_per_step"], (d.get("e2e") or {}).get("value"), d["loss"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/b2.err").read()[-1500:])
PY
