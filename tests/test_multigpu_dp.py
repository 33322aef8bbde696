"""2+ GPU data-parallel test: fused all-reduce+GD kernel vs NCCL baseline vs single-GPU trajectory."""
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("model", ["cifar_convnet", "mnist_mlp", "wide_mlp_ae", "lstm_classifier"])
def test_dp_matches_single_gpu(model):
    import torch
    n = min(torch.cuda.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n),
           "--master-addr", "127.0.0.1", "--master-port", "29611", os.path.join(ROOT, "scripts", "dp_check.py"), model]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600, cwd=ROOT)
    out = r.stdout.decode(errors="replace")
    assert r.returncode == 0 and "DP_CHECK OK" in out, out[-4000:]
