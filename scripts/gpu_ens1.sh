S=veles_b200/znicz/samples/cifar.py
T0=$(date +%s.%N)
timeout 600 python -m veles_b200 --ensemble-train 8:0.9 --result-file gpurun_out/ens1.json $S - root.cifar.decision.max_epochs=3 root.cifar.loader.n_train=50000 root.cifar.loader.n_valid=5000 root.common.disable.snapshotting=False "root.common.dirs.snapshots=\"gpurun_out/ens_snaps\"" > gpurun_out/ens1.log 2>&1; echo ens1 rc=$?
T1=$(date +%s.%N); echo "ensemble of 8 on 1 GPU: $(python -c "print(round($T1-$T0,1))") s"; tail -3 gpurun_out/ens1.log | cut -c1-300; rm -rf gpurun_out/ens_snaps
python -c "import json; d=json.load(open('gpurun_out/ens1.json')); print(len(d['models']), [round(m.get('EvaluationFitness',0),3) for m in d['models']])"
