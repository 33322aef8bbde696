#!/bin/bash
# 2 GPUs: device-side class-end reductions + gather (dp_check), then the CLI as a data-parallel job vs one GPU
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29721 scripts/dp_check.py cifar_convnet > gpurun_out/dp_check_cifar2.log 2>&1; echo dpcheck rc=$?; grep -E "DP_CHECK|differ|Error|error" gpurun_out/dp_check_cifar2.log | cut -c1-200 | head -12
S=veles_b200/znicz/samples/mnist.py
ARGS="root.mnist.decision.max_epochs=3 root.mnist.loader.n_train=6000 root.mnist.loader.n_valid=1000 root.common.disable.snapshotting=True"
timeout 200 $TR --master-port 29722 -m veles_b200 --no-logo -v warning -r 1234 --result-file gpurun_out/cli_dp2.json $S - $ARGS > gpurun_out/cli_dp2.log 2>&1; echo cli-dp2 rc=$?; tail -3 gpurun_out/cli_dp2.log | cut -c1-300
timeout 200 python -m veles_b200 --no-logo -v warning -r 1234 --result-file gpurun_out/cli_dp1.json $S - $ARGS > gpurun_out/cli_dp1.log 2>&1; echo cli-1gpu rc=$?; tail -3 gpurun_out/cli_dp1.log | cut -c1-300
python - <<'PY'
import json
for f in ("gpurun_out/cli_dp2.json", "gpurun_out/cli_dp1.json"):
    try:
        d = json.load(open(f))
        cm = d.get("Confusion", {}).get("validation") or d.get("Confusion", {}).get("train")
        print(f, {k: d[k] for k in d if k in ("EvaluationFitness", "Best epoch", "Total epochs", "Errors")}, "confusion sum", sum(map(sum, cm)) if cm else None)
    except Exception as e:
        print(f, "unreadable", e)
PY
