"""Data-parallel correctness check, run under torchrun with N >= 2 ranks:
N-rank training with the fused all-reduce+GD kernel (and with the NCCL baseline) must follow the
same weight trajectory as 1-rank training on the same GLOBAL minibatches."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist


def run(model, B, steps, dp, rank, world, device, layers_kwargs=None, use_graph=True):
    from veles_b200.trainer import Trainer
    tr = Trainer(model, minibatch_size=B, device=device, use_graph=use_graph, data_parallel=dp, n_train=4 * B,
                 layers_kwargs=layers_kwargs)
    for _ in range(steps):
        tr.step_resident()
    torch.cuda.synchronize()
    bias_same = True
    if dp is not None:
        dp.check()
        if dp.mode == "fused":
            # BEFORE any gather: what the forward passes read (bias = fp32 master, bf16 weight shadow) must already be
            # identical on every rank after several steps - a value that only its owner holds would show up here
            for f in tr.wf.forwards:
                if not getattr(f, "HAS_WEIGHTS", False):
                    continue
                tensors = [f.bias.devmem] if (f.bias is not None and f.bias) else []
                if dp.weight_delivery != "peer_tma" and getattr(f, "weights_bf16_", None) is not None:
                    tensors.append(f.weights_bf16_)
                for t in tensors:
                    t0 = t.clone()
                    dist.broadcast(t0, 0)
                    if not torch.equal(t, t0):
                        bias_same = False
                        print("rank %d: %s of %s differs from rank 0 before sync_master" % (
                            rank, "bias" if t.dtype == torch.float32 else "shadow", f.name), flush=True)
        dp.sync_master()
        if not counters_ok(dp, rank, world):
            bias_same = False
        if dp.mode == "fused" and dp.weight_delivery == "peer_tma":
            for f in tr.wf.forwards:   # non-owned shadow rows are stale by design: rebuild them for the comparison
                if getattr(f, "peer_weights_", None) is not None:
                    f.refresh_shadow()
            torch.cuda.synchronize()
    out = []
    gd_of = {id(g.forward): g for g in tr.wf.gds if getattr(g, "forward", None) is not None}
    for f in tr.wf.forwards:
        if getattr(f, "HAS_WEIGHTS", False):
            f.weights.map_read()
            f.bias.map_read()
            vel = None
            g = gd_of.get(id(f))
            if g is not None and g.gradient_weights_with_moment:
                g.gradient_weights_with_moment.map_read()
                vel = g.gradient_weights_with_moment.mem.copy()
            out.append((f.weights.mem.copy(), f.bias.mem.copy(), f.weights_bf16_.float().cpu().numpy().copy(), vel))
    tr.close()
    if not bias_same:
        out.append("MISMATCH")
    return out


def counters_ok(dp, rank, world):
    """The class-end reductions of the group (device-side exchange in fused mode, the process group otherwise):
    counters + loss, max / min, a confusion matrix - several rounds, so both slot parities are reused."""
    ok = True
    ranks = np.arange(1, world + 1)
    for it in range(5):
        vals, loss = dp.sum_counters([3 * (rank + 1) + it, 100 + rank], 0.125 * (rank + 1))
        ok &= list(vals) == [int(3 * ranks.sum() + it * world), int(100 * world + ranks.sum() - world)]
        ok &= abs(loss - 0.125 * ranks.sum()) < 1e-12
        mx, mn = dp.max_min(float(rank) + it, -2.0 * rank)
        ok &= (mx, mn) == (float(world - 1 + it), -2.0 * (world - 1))
        cm = dp.sum_array((np.arange(100, dtype=np.int32).reshape(10, 10) + it) * (rank + 1))
        ok &= cm.dtype == np.int32 and np.array_equal(cm, (np.arange(100).reshape(10, 10) + it) * ranks.sum())
        top = dp.sum_array(np.array([0.5 * rank], dtype=np.float32), op="max")
        ok &= float(top[0]) == 0.5 * (world - 1)
    if not ok:
        print("rank %d: class-end reductions differ from the expected sums (mode %s)" % (rank, dp.mode), flush=True)
    return bool(ok)


model = "cifar_convnet"


def main():
    global model
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    rank, world = dist.get_rank(), dist.get_world_size()
    from veles_b200.backends import B200Device
    from veles_b200.parallel.dp import DataParallel
    device = B200Device(index=local_rank)
    model = sys.argv[1] if len(sys.argv) > 1 else "cifar_convnet"
    B, steps = (16, 4) if model == "alexnet" else (32, 6)
    kw = {"caffe_init": False, "lr": 0.01} if model == "cifar_convnet" else None
    if model == "alexnet":
        kw = {"dropout": 0.0}     # dropout masks depend on the per-rank batch: no common trajectory with them on
    results = {}
    modes = ("fused", "nccl") if model == "alexnet" else ("fused", "fused_pull", "fused_peer_tma", "nccl")
    if len(sys.argv) > 2:
        modes = tuple(sys.argv[2].split(","))
    for mode in modes:
        for graph in ((True, False) if mode == "fused" else (False,) if mode == "nccl" else (True,)):
            if mode == "fused_peer_tma":
                dp = DataParallel(mode="fused", weight_delivery="peer_tma")
            elif mode == "fused_pull":
                dp = DataParallel(mode="fused", exchange="pull")
            else:
                dp = DataParallel(mode=mode)
            results[(mode, graph)] = run(model, B, steps, dp, rank, world, device, kw, use_graph=graph)
            if rank == 0:
                n_peer = sum(1 for i in dp.unit_info.values() if i.get("peer_shadow")) if mode != "nccl" else 0
                print("mode=%s graph=%s multimem=%s world=%d peer-TMA layers=%d" % (mode, graph, dp.use_multimem, world, n_peer),
                      flush=True)
            dist.barrier()
    ok = True
    if rank == 0:
        # single-GPU reference on the same global minibatches (global batch = world * B)
        ref = run(model, B * world, steps, None, 0, 1, device, kw, use_graph=False)
        for key, res in results.items():
            if res and isinstance(res[-1], str):
                ok = False
                res.pop()
            for li, ((w, b, sh, vel), (rw, rb, rsh, rvel)) in enumerate(zip(res, ref)):
                scale = np.abs(rw).max() + 1e-9
                ew = np.abs(w - rw).max() / scale
                eb = np.abs(b - rb).max() / max(np.abs(rb).max(), 1e-3 * scale, 1e-6)
                es = np.abs(sh - w).max() / scale
                ev = 0.0
                if vel is not None and rvel is not None and key[0] != "nccl":
                    ev = np.abs(vel - rvel).max() / (np.abs(rvel).max() + 1e-9)   # gathered from the owners
                print("mode=%s graph=%s layer %d: |dW|/|W| = %.3e  |db| = %.3e  shadow-vs-master = %.3e  |dvel| = %.3e"
                      % (key[0], key[1], li, ew, eb, es, ev), flush=True)
                # AlexNet-sized runs: bf16 activations re-rounded through eight layers make the tiny gradients of a
                # random-init net differ by a few per cent between batch splits (the NCCL arm shows the same deviations), so the
                # strict bounds apply to the small deterministic models and a loose one to the big one
                tol_b, tol_v = (2e-1, 4e-1) if model == "alexnet" else (5e-2, 5e-2)
                if not (ew < 2e-2 and eb < tol_b and es < 1e-2 and ev < tol_v and np.isfinite(w).all()):
                    ok = False
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    # every rank must hold identical weights
    for key, res in results.items():
        if res and isinstance(res[-1], str):
            res.pop()
            flag.zero_()
        for w, b, sh, vel in res:
            t = torch.from_numpy(w).cuda()
            t0 = t.clone()
            dist.broadcast(t0, 0)
            if not torch.equal(t, t0):
                print("rank %d: weights differ from rank 0 in mode %s" % (rank, key), flush=True)
                flag.zero_()
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:   # one writer: concurrent prints of several ranks interleave
        sys.stdout.write("DP_CHECK %s\n" % ("OK" if int(flag.item()) else "FAILED"))
        sys.stdout.flush()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
