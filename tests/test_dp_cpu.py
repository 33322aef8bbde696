"""Data parallelism without GPUs: two gloo processes on the numpy backend (DataParallel(mode="host")) must follow
the single-process trajectory on the same global minibatches - the CPU oracle of the fused all-reduce + update kernel
and the test of the host-side logic (rank-sharded loader, gradient reduction, identical replicas)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("model,port", [("mnist_mlp", 29661), ("cifar_convnet", 29662), ("wide_mlp_ae", 29663),
                                        ("lstm_classifier", 29664)])
def test_two_gloo_ranks_match_single_process(model, port):
    env = dict(os.environ, OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "scripts", "dp_check_cpu.py"), model]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600, cwd=ROOT, env=env)
    out = r.stdout.decode(errors="replace")
    assert r.returncode == 0 and "DP_CHECK_CPU OK" in out, out[-3000:]
