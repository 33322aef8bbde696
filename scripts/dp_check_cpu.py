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
    dp = DataParallel(mode="host")
    res = run(model, B, steps, dp, kw)
    ok = True
    # the class-end reductions of the group (what Decision calls): counters + loss, max / min, small arrays
    ranks = np.arange(1, world + 1)
    vals, loss = dp.sum_counters([3 * (rank + 1), 100 + rank], 0.125 * (rank + 1))
    ok = ok and list(vals) == [int(3 * ranks.sum()), int(100 * world + ranks.sum() - world)]
    ok = ok and abs(loss - 0.125 * ranks.sum()) < 1e-12
    ok = ok and dp.max_min(float(rank), -2.0 * rank) == (float(world - 1), -2.0 * (world - 1))
    cm = dp.sum_array(np.arange(12, dtype=np.int32).reshape(3, 4) * (rank + 1))
    ok = ok and cm.dtype == np.int32 and np.array_equal(cm, np.arange(12).reshape(3, 4) * ranks.sum())
    ok = ok and float(dp.sum_array(np.array([0.5 * rank], dtype=np.float32), op="max")[0]) == 0.5 * (world - 1)
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
