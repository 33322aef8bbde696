"""CPU (gloo) data-parallel check: N processes on the numpy backend with DataParallel(mode="host") must follow the
weight trajectory of one process training on the same GLOBAL minibatches.  Run under torchrun:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29650 \\
        scripts/dp_check_cpu.py [model]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist


def run(model, B, steps, dp, kw):
    from veles_b200.backends import NumpyDevice
    from veles_b200.trainer import Trainer
    tr = Trainer(model, minibatch_size=B, device=NumpyDevice(), use_graph=False, data_parallel=dp, n_train=4 * B,
                 layers_kwargs=kw)
    for _ in range(steps):
        tr.step_resident()
    out = [(f.weights.mem.copy(), f.bias.mem.copy()) for f in tr.wf.forwards if getattr(f, "HAS_WEIGHTS", False)]
    tr.close()
    return out


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    from veles_b200.parallel.dp import DataParallel
    model = sys.argv[1] if len(sys.argv) > 1 else "mnist_mlp"
    kw = {"caffe_init": False, "lr": 0.01} if model == "cifar_convnet" else None
    B, steps = 16, 4
    res = run(model, B, steps, DataParallel(mode="host"), kw)
    ok = True
    if rank == 0:
        ref = run(model, B * world, steps, None, kw)
        for li, ((w, b), (rw, rb)) in enumerate(zip(res, ref)):
            ew = np.abs(w - rw).max() / (np.abs(rw).max() + 1e-9)
            eb = np.abs(b - rb).max() / (np.abs(rb).max() + 1e-6)
            print("layer %d: |dW|/|W| = %.3e  |db| = %.3e" % (li, ew, eb), flush=True)
            ok = ok and ew < 1e-4 and eb < 1e-3 and np.isfinite(w).all()
    flag = torch.tensor([1 if ok else 0])
    dist.broadcast(flag, 0)
    for w, b in res:   # every rank holds the same weights
        t = torch.from_numpy(w.copy())
        t0 = t.clone()
        dist.broadcast(t0, 0)
        if not torch.equal(t, t0):
            flag.zero_()
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:   # one writer: concurrent prints of several ranks interleave
        sys.stdout.write("DP_CHECK_CPU %s\n" % ("OK" if int(flag.item()) else "FAILED"))
        sys.stdout.flush()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
