#!/usr/bin/env python
"""Headline benchmark: AlexNet (ImageNet shape 227x227x3, 1000 classes) training throughput in
images/s on N B200s of one node (BASELINE.json: "CIFAR-10/AlexNet images/sec, 1/2/4/8 B200").

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Prints ONE JSON line on rank 0.  ``value`` is the whole-job images/s of the device-timed
training step (forward + evaluator + backward + fused GD update [+ fused gradient all-reduce]);
``e2e`` is the same measured through the public ``veles_b200.Trainer.step`` API including the
per-step pinned-host -> device upload of the minibatch and the device -> host read of the loss.
Weak scaling: the per-GPU minibatch is fixed, the global batch grows with N.

``--impl reference`` would run the unmodified reference from baseline/_ref; it cannot be
installed here (see DESIGN.md "Reference arm"), so it prints an "unavailable" line.
``--impl torch|nccl|ps_emulation`` are stand-in comparison arms (baseline/standins.py).
``--model cifar_convnet|wide_mlp_ae|mnist_mlp`` benchmarks the other BASELINE configs.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="veles_b200", choices=["veles_b200", "reference", "nccl", "torch", "ps_emulation"],
                   help="veles_b200 (the product); reference (unavailable here); stand-ins: nccl = this engine with an "
                        "NCCL all-reduce + separate GD kernel, torch = stock torch.nn/cuDNN/DDP, ps_emulation = the "
                        "reference's ZeroMQ master-slave schedule around torch compute (baseline/standins.py)")
    p.add_argument("--min-seconds", type=float, default=0.0,
                   help="repeat the timed region (in whole multiples of --steps) until it has run this long: sustained "
                        "clocks instead of a burst")
    p.add_argument("--exchange", default="pull", choices=["push", "pull"],
                   help="N>1: gradients reach their owner by remote stores (push) or by multimem.ld_reduce / peer loads (pull)")
    p.add_argument("--no-multimem", action="store_true", help="N>1: peer loads / stores instead of NVLS multimem")
    p.add_argument("--dp-profile", action="store_true",
                   help="N>1, with --no-graph: print the duration of every exchange kernel of one step to stderr")
    p.add_argument("--bucket-mb", type=float, default=4.0, help="N>1: parameter bytes that close an exchange bucket")
    p.add_argument("--fc-split-k", action="store_true",
                   help="small-batch fully connected layers: split-K GEMM into an fp32 scratch + bias/activation pass")
    p.add_argument("--pdl", type=int, default=-1, choices=[-1, 0, 1],
                   help="programmatic dependent launch of this library's kernels (-1: the engine default)")
    p.add_argument("--e2e-sync", action="store_true",
                   help="end-to-end region: wait for each step's loss before enqueueing the next step (default: the "
                        "loss is read back one step late, every step)")
    p.add_argument("--ps-async", action="store_true", help="ps_emulation: two jobs in flight per slave (--async-slave)")
    p.add_argument("--model", default="alexnet")
    p.add_argument("--batch", type=int, default=0, help="per-GPU minibatch (default: model's)")
    p.add_argument("--no-graph", action="store_true")
    p.add_argument("--skip-e2e", action="store_true")
    p.add_argument("--weight-delivery", default="multicast", choices=["multicast", "peer_tma"],
                   help="N>1: multimem.st broadcast of updated bf16 weights, or GEMMs TMA-loading weight tiles from owners")
    p.add_argument("--comm-ctas", type=int, default=0,
                   help="N>1: CTAs of the fused all-reduce+GD kernel = SMs the backward GEMMs leave free for it "
                        "(0 = by world size: 16 up to 4 ranks, else 8)")
    p.add_argument("--gemm-schedule", default="static", choices=["static", "clc"],
                   help="GEMM tile scheduling: static persistent round-robin, or cluster-launch-control work stealing")
    p.add_argument("--split-handshake", action="store_true",
                   help="N>1: await the peers' ready flags in a separate 1-CTA launch ahead of the worker CTAs")
    p.add_argument("--replicate-master", action="store_true", help="N>1: also broadcast the fp32 master weights each step")
    return p.parse_args()


def reference_unavailable():
    why = ("devbib/veles cannot be installed/run on this image: Znicz NN engine submodule is absent from the "
           "checkout, core uses Python<3.7-only syntax (async attribute), deps twisted/zope/cuda4py not in wheelhouse")
    marker = os.path.join(ROOT, "baseline", "INSTALL_OUTCOME.txt")
    if os.path.exists(marker):
        with open(marker) as f:
            first = f.readline().strip()
        if first:
            why = first
    print(json.dumps({"impl": "reference", "unavailable": why}))


def _pdl_state():
    from veles_b200.ops import kernels as _k
    on = _k.set_pdl(False)
    _k.set_pdl(on)
    return bool(on)


def main():
    from veles_b200.config import root as _cfg
    _cfg.common.disable.spinning_run_progress = True      # stdout carries exactly one JSON line
    args = parse()
    if args.impl == "reference":
        reference_unavailable()
        return 0
    import torch
    import torch.distributed as dist
    from veles_b200.models.configs import MODEL_CONFIGS
    from veles_b200.ops import kernels
    from veles_b200.utils.clocks import ClockSampler

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world:
        if world == 1 and args.gpus > 1:
            # allow `python bench.py --gpus N` without torchrun: re-exec under torch.distributed.run
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
                   "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 2000)] + sys.argv
            os.execv(sys.executable, cmd)
        raise SystemExit("--gpus %d does not match WORLD_SIZE %d" % (args.gpus, world))
    torch.cuda.set_device(local_rank)
    dp = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if args.impl in ("torch", "ps_emulation"):
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import standins
        args.batch = args.batch or MODEL_CONFIGS[args.model]["minibatch"]

        def _sync():
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
                torch.cuda.synchronize()

        def _max(ms):
            if world == 1:
                return ms
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())

        fn = standins.run_torch if args.impl == "torch" else standins.run_ps_emulation
        fn(args, world, rank, local_rank, _sync, _max, ClockSampler)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0
    if world > 1:
        from veles_b200.parallel.dp import DataParallel
        dp = DataParallel(mode="nccl" if args.impl == "nccl" else "fused", weight_delivery=args.weight_delivery,
                          shard_master=not args.replicate_master, n_ctas=args.comm_ctas,
                          split_handshake=args.split_handshake, bucket_bytes=int(args.bucket_mb * (1 << 20)),
                          exchange=args.exchange, use_multimem=False if args.no_multimem else None)

    from veles_b200.backends import B200Device
    from veles_b200.trainer import Trainer
    if args.gemm_schedule == "clc":
        from veles_b200.ops import kernels as _k
        _k.set_dynamic_schedule(True)
    if args.fc_split_k:
        from veles_b200.config import root as _root
        _root.common.engine.two_phase_split_k = True
    if args.pdl >= 0:
        from veles_b200.ops import kernels as _k
        _k.set_pdl(bool(args.pdl))
    cfg = MODEL_CONFIGS[args.model]
    B = args.batch or cfg["minibatch"]
    device = B200Device(index=local_rank)
    use_graph = not args.no_graph and args.impl != "nccl"

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------------------------------------------------------- kernel-only (device resident data)
    trainer = Trainer(args.model, minibatch_size=B, device=device, use_graph=use_graph, data_parallel=dp,
                      n_train=8 * B)
    for _ in range(max(args.warmup, 3)):
        trainer.step_resident()
    # count launches of one step eagerly (graph replays do not pass through the Python wrappers)
    sync_all()
    trainer.stepper.use_graph = False
    l0 = kernels.LAUNCHES[0]
    trainer.step_resident()
    launches_per_step = kernels.LAUNCHES[0] - l0
    trainer.stepper.use_graph = use_graph
    for _ in range(3):
        trainer.step_resident()
    sync_all()
    if args.dp_profile and dp is not None and not use_graph and args.impl != "nccl":
        dp.profile = True
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        trainer.step_resident()
        e1.record()
        times = dp.bucket_times()
        dp.profile = False
        if rank == 0:
            sys.stderr.write("exchange kernels of one eager step (%.3f ms): %s\n" % (
                e0.elapsed_time(e1), ", ".join("bucket %d %.1f MB %.3f ms" % (c, b / 1e6, t) for c, b, t in times)))
        sync_all()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    # L2 policy: AlexNet streams far more than the 126 MB L2 per step; the small models would otherwise re-hit
    # their previous step's lines, so L2 is flushed (a 256 MB write) between steps and every step is timed on its own.
    flush_buf = None if args.model == "alexnet" else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    steps_timed = args.steps
    trainer.read_loss()     # the evaluator's loss sum restarts here: `loss` below is the mean over the timed steps
    if flush_buf is None:
        ms_total, steps_timed = 0.0, 0
        while True:
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.steps):
                trainer.step_resident()
            e1.record()
            sync_all()
            ms_total += max_over_ranks(e0.elapsed_time(e1))
            steps_timed += args.steps
            if ms_total >= args.min_seconds * 1000.0:
                break
    else:
        evs = []
        for _ in range(args.steps):
            flush_buf.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            trainer.step_resident()
            b.record()
            evs.append((a, b))
        sync_all()
        ms_total = max_over_ranks(sum(a.elapsed_time(b) for a, b in evs))
    clocks = sampler.stop() if sampler is not None else None
    loss_after = trainer.read_loss() / steps_timed
    ms_per_step = ms_total / steps_timed
    value = world * B * 1000.0 / ms_per_step
    trainer.close()
    del trainer
    torch.cuda.empty_cache()

    # ---------------------------------------------------------------- end to end (public API, host-fed)
    e2e = None
    if not args.skip_e2e:
        import numpy as np
        dp2 = None
        if world > 1:
            from veles_b200.parallel.dp import DataParallel
            dp2 = DataParallel(mode="nccl" if args.impl == "nccl" else "fused", weight_delivery=args.weight_delivery,
                               shard_master=not args.replicate_master, n_ctas=args.comm_ctas,
                               split_handshake=args.split_handshake, bucket_bytes=int(args.bucket_mb * (1 << 20)),
                          exchange=args.exchange, use_multimem=False if args.no_multimem else None)
        tr = Trainer(args.model, minibatch_size=B, device=device, use_graph=use_graph, streaming=True,
                     data_parallel=dp2)
        shape = cfg["sample_shape"]
        image_like = len(shape) == 3
        rng = np.random.RandomState(100 + rank)
        pool = []
        for i in range(4):
            if image_like:
                d = torch.from_numpy(rng.randint(0, 256, size=(B,) + shape).astype(np.uint8)).pin_memory()
            else:
                d = torch.from_numpy(rng.uniform(-1, 1, size=(B,) + shape).astype(np.float32)).pin_memory()
            lab = torch.from_numpy(rng.randint(0, max(cfg["n_classes"], 2), size=B).astype(np.int32)).pin_memory()
            pool.append((d, lab))
        h2d = 0
        # double buffering: minibatch i+1 is uploaded while minibatch i trains
        tr.feed(*pool[0])
        for i in range(max(args.warmup, 3) + 3):
            tr.feed(*pool[(i + 1) % 4])
            tr.step(prefetched=True)
        def timed_region(sync_every_step):
            h2d_bytes, loss = 0, None
            sync_all()
            t0 = time.perf_counter()
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            if sync_every_step:
                # the host waits for every step's loss before it enqueues the next step
                for i in range(args.steps):
                    h2d_bytes = tr.feed(*pool[(i + 1) % 4])
                    loss = tr.step(prefetched=True)
            else:
                # every step's loss still crosses to the host (4 bytes, pinned), but it is read one step late: the
                # host enqueues step i+1 while the device runs step i
                for i in range(args.steps):
                    h2d_bytes = tr.feed(*pool[(i + 1) % 4])
                    prev = tr.step_pipelined(prefetched=True)
                    if prev is not None:
                        loss = prev
                loss = tr.flush_loss()
            e1.record()
            sync_all()
            wall = (time.perf_counter() - t0) * 1000.0
            return loss, h2d_bytes, max_over_ranks(e0.elapsed_time(e1)), wall

        e2e_sync = bool(args.e2e_sync)
        try:
            last_loss, h2d, dev_ms, wall_ms = timed_region(e2e_sync)
        except Exception as exc:  # noqa: BLE001 - the late-read path must never cost the record
            if e2e_sync:
                raise
            sys.stderr.write("pipelined loss read-back failed (%s): timing the synchronous form\n" % exc)
            e2e_sync = True
            while len(getattr(tr.wf.loader, "pending_", ())) >= 2:     # leave room for the region's first upload
                tr.step(prefetched=True)
            last_loss, h2d, dev_ms, wall_ms = timed_region(True)
        e2e_ms = max(dev_ms, max_over_ranks(wall_ms)) / args.steps
        e2e = {"value": world * B * 1000.0 / e2e_ms, "unit": "images/s", "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4, "last_loss": last_loss,
               "api": ("veles_b200.Trainer.feed/step (pinned host uint8 -> device, loss read back each step)"
                       if e2e_sync else
                       "veles_b200.Trainer.feed/step_pipelined (pinned host uint8 -> device every step; every step's "
                       "loss copied to pinned host memory and read one step late, the last one before the clock stops)")}
        tr.close()

    if rank == 0:
        # BASELINE.md publishes no throughput; the denominators are the stand-in arms measured on this pool
        # (baseline/standin_numbers.json, written from profiles/bench records): stock torch.nn/cuDNN/DDP first
        vs_baseline, vs_source = None, None
        try:
            with open(os.path.join(ROOT, "baseline", "standin_numbers.json")) as f:
                rec = json.load(f).get(args.model, {}).get(str(world), {})
            for arm in ("torch", "nccl", "ps_emulation"):
                if rec.get(arm):
                    vs_baseline = value / float(rec[arm])
                    vs_source = "stand-in '%s' at %d GPU(s): %.0f images/s (not the reference; BASELINE.md section 3)" % (
                        arm, world, float(rec[arm]))
                    break
        except (OSError, ValueError):
            pass
        out = {
            "metric": "AlexNet images/sec" if args.model == "alexnet" else "%s samples/sec" % args.model,
            "value": value, "unit": "images/s", "n_gpus": world, "steps": steps_timed, "warmup": max(args.warmup, 3) + 4,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": vs_baseline,
            "vs_baseline_source": vs_source,
            "dtype": "bf16", "data": "synthetic",
            "config": {"model": args.model, "global_batch": world * B, "per_gpu_batch": B,
                       "input_shape": list(cfg["sample_shape"]),
                       "parallelism": "dp%d" % world if world > 1 else "single",
                       "grad_exchange": ("none" if world == 1 else
                                         ("nccl_allreduce+gd" if args.impl == "nccl" else
                                          ("fused_push_reduce_gd_multimem_bcast" if args.exchange == "push"
                                           else "fused_multimem_allreduce_gd"))),
                       "weight_delivery": (None if world == 1 or args.impl == "nccl" else args.weight_delivery),
                       "master_weights": (None if world == 1 or args.impl == "nccl" else
                                          ("replicated" if args.replicate_master else "sharded, gathered at epoch ends")),
                       "gemm_schedule": args.gemm_schedule,
                       "programmatic_dependent_launch": _pdl_state(),
                       "cuda_graph": use_graph,
                       "l2_policy": ("inputs larger than L2: every step gathers a new minibatch from the resident "
                                     "dataset and streams activations >> 126 MB" if args.model == "alexnet" else
                                     "L2 flushed (256 MB write) between steps; each step timed with its own CUDA events"),
                       "precision": "bf16 operands/activations, fp32 accumulate, fp32 master weights"},
            "impl": args.impl, "gpu_launches": int(launches_per_step * steps_timed),
            "launches_per_step": int(launches_per_step), "clocks": clocks, "e2e": e2e, "loss": loss_after,
        }
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
