# 8-GPU records of round 2 (one gpurun --gpus 8 call): AlexNet at 8 and 4 ranks, stand-ins, configs 2 / 4 / 5.
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
show() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d.get("impl"), d.get("n_gpus"), round(d["value"]), round(d["ms_per_step"], 3), (round(d["e2e"]["value"]) if d.get("e2e") else None))
except Exception as e:
    print(sys.argv[1], "NO RECORD", e)
PY
}
run() { # name timeout cmd...
  n=$1; t=$2; shift 2
  timeout $t "$@" > gpurun_out/$n.json 2> gpurun_out/$n.err; echo "$n rc=$?"; show gpurun_out/$n.json
}
run a8_c8   200 $TR8 --master-port 29801 bench.py --gpus 8 --steps 20 --warmup 5 --comm-ctas 8
run a8_c16  200 $TR8 --master-port 29802 bench.py --gpus 8 --steps 20 --warmup 5 --skip-e2e --comm-ctas 16
# two 4-rank jobs side by side on disjoint GPUs
(CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 200 $TR4 --master-port 29803 bench.py --gpus 4 --steps 20 --warmup 5 --skip-e2e --comm-ctas 16 > gpurun_out/a4_c16.json 2> gpurun_out/a4_c16.err) &
(CUDA_VISIBLE_DEVICES=4,5,6,7 timeout 200 $TR4 --master-port 29804 bench.py --gpus 4 --steps 20 --warmup 5 --skip-e2e --comm-ctas 8 > gpurun_out/a4_c8.json 2> gpurun_out/a4_c8.err) &
wait; show gpurun_out/a4_c16.json; show gpurun_out/a4_c8.json
run a8_nccl  200 $TR8 --master-port 29805 bench.py --gpus 8 --steps 20 --warmup 5 --skip-e2e --impl nccl
run a8_torch 240 $TR8 --master-port 29806 bench.py --gpus 8 --steps 20 --warmup 5 --skip-e2e --impl torch
run cifar8   200 $TR8 --master-port 29807 bench.py --gpus 8 --steps 30 --warmup 5 --skip-e2e --model cifar_convnet
run mlp8     200 $TR8 --master-port 29809 bench.py --gpus 8 --steps 30 --warmup 5 --skip-e2e --model wide_mlp_ae
run mlp8_nccl 200 $TR8 --master-port 29810 bench.py --gpus 8 --steps 30 --warmup 5 --skip-e2e --model wide_mlp_ae --impl nccl
timeout 240 $TR8 --master-port 29811 scripts/dp_check.py cifar_convnet fused > gpurun_out/dp_check_cifar8.log 2>&1; echo dpcheck8 rc=$?; grep -E "DP_CHECK|differs" gpurun_out/dp_check_cifar8.log | head -5
# config 5: eight CIFAR models, one per GPU (wall clock of the whole ensemble run)
S=veles_b200/znicz/samples/cifar.py
T0=$(date +%s.%N)
timeout 400 python -m veles_b200 --ensemble-train 8:0.9 --result-file gpurun_out/ens8.json $S - root.cifar.decision.max_epochs=3 root.cifar.loader.n_train=50000 root.cifar.loader.n_valid=5000 root.common.disable.snapshotting=False "root.common.dirs.snapshots=\"gpurun_out/ens_snaps\"" > gpurun_out/ens8.log 2>&1; echo ens8 rc=$?
T1=$(date +%s.%N); echo "ensemble of 8 on 8 GPUs: $(python -c "print(round($T1-$T0,1))") s"; tail -2 gpurun_out/ens8.log | cut -c1-200; rm -rf gpurun_out/ens_snaps
