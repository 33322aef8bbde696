#!/bin/bash
# A/B of programmatic dependent launch: tests with it on, then bench pairs
mkdir -p gpurun_out
VELES_B200_PDL=1 timeout 400 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_workflow.py tests/test_gpu_dp_emulated.py -x -q > gpurun_out/pdl_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/pdl_tests.log
for m in alexnet cifar_convnet wide_mlp_ae; do
  for p in 0 1; do
    timeout 200 python bench.py --model $m --steps 30 --warmup 5 --skip-e2e --pdl $p > gpurun_out/pdl_${m}_$p.json 2> gpurun_out/pdl_${m}_$p.err; echo "$m pdl=$p rc=$?"
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/pdl_${m}_$p.json").read().strip().splitlines()[-1]); print("  ", d["value"], d["ms_per_step"], d.get("launches_per_step"), d.get("loss"))
except Exception as e:
    print("  parse failed", e); print(open("gpurun_out/pdl_${m}_$p.err").read()[-1500:])
PY
  done
done
