TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29702 scripts/dp_check.py alexnet fused,nccl > gpurun_out/dp_check_alexnet.log 2>&1; echo dpcheck-alexnet rc=$?; grep -E "DP_CHECK|differs|layer|Error" gpurun_out/dp_check_alexnet.log | cut -c1-170 | head -30
for c in 8 16 24 32; do
timeout 200 $TR --master-port 297$c bench.py --gpus 2 --steps 20 --warmup 5 --skip-e2e --comm-ctas $c > gpurun_out/bench_2gpu_pull_c$c.json 2> gpurun_out/bench_2gpu_pull_c$c.err; echo pull ctas $c rc=$?; python -c "import json; d=json.loads(open('gpurun_out/bench_2gpu_pull_c$c.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'])"
done
timeout 200 $TR --master-port 29750 bench.py --gpus 2 --steps 20 --warmup 5 --skip-e2e --comm-ctas 16 --no-multimem > gpurun_out/bench_2gpu_p2p_c16.json 2> gpurun_out/bench_2gpu_p2p_c16.err; echo p2p ctas 16 rc=$?; python -c "import json; d=json.loads(open('gpurun_out/bench_2gpu_p2p_c16.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'])"
