"""Stand-in comparison arms for ``bench.py`` (NOT the reference: the reference cannot run here).

``--impl torch``         stock ``torch.nn`` model of the same architecture (cuDNN / cuBLAS kernels, bf16 autocast,
                         channels_last, fp32 master weights, ``torch.optim.SGD`` with momentum and weight decay,
                         ``DistributedDataParallel`` over NCCL for N > 1): what a user gets from the library stack.
``--impl ps_emulation``  the reference's master-slave schedule with the same torch compute on every slave: per
                         minibatch the slave receives the pickled weights of a job over ZeroMQ (ipc), uploads them,
                         trains one minibatch, downloads the weight delta (``Array.map_read``), pickles it into an
                         "update", and the master adds it under a lock and pickles the next job
                         (``/root/reference/veles/server.py:141-169, 401-430, 596-611``,
                         ``veles/client.py:138-160, 278-342``, ``veles/workflow.py:456-573``); at most two jobs in
                         flight per slave (``--async-slave`` behaviour, ``veles/client.py:294``).

Both print the same JSON line as the main arm, with ``"impl"`` naming the stand-in and ``"stand_in": true``.
None of veles_b200's models, kernels or engine is on these paths.
"""
from __future__ import annotations

import json
import os
import pickle
import threading
import time


def build_torch_model(name, n_classes=None):
    import torch.nn as nn
    if name == "alexnet":
        n_classes = n_classes or 1000
        return nn.Sequential(
            nn.Conv2d(3, 96, 11, stride=4), nn.ReLU(inplace=True), nn.LocalResponseNorm(5, 1e-4, 0.75, 1.0),
            nn.MaxPool2d(3, 2),
            nn.Conv2d(96, 256, 5, padding=2, groups=2), nn.ReLU(inplace=True), nn.LocalResponseNorm(5, 1e-4, 0.75, 1.0),
            nn.MaxPool2d(3, 2),
            nn.Conv2d(256, 384, 3, padding=1), nn.ReLU(inplace=True),
            nn.Conv2d(384, 384, 3, padding=1, groups=2), nn.ReLU(inplace=True),
            nn.Conv2d(384, 256, 3, padding=1, groups=2), nn.ReLU(inplace=True), nn.MaxPool2d(3, 2),
            nn.Flatten(), nn.Linear(256 * 6 * 6, 4096), nn.ReLU(inplace=True), nn.Dropout(0.5),
            nn.Linear(4096, 4096), nn.ReLU(inplace=True), nn.Dropout(0.5), nn.Linear(4096, n_classes))
    if name == "cifar_convnet":
        n_classes = n_classes or 10
        return nn.Sequential(
            nn.Conv2d(3, 32, 5, padding=2), nn.ReLU(inplace=True), nn.MaxPool2d(3, 2, ceil_mode=True),
            nn.Conv2d(32, 32, 5, padding=2), nn.ReLU(inplace=True), nn.AvgPool2d(3, 2, ceil_mode=True),
            nn.Conv2d(32, 64, 5, padding=2), nn.ReLU(inplace=True), nn.AvgPool2d(3, 2, ceil_mode=True),
            nn.Flatten(), nn.Linear(64 * 4 * 4, 64), nn.ReLU(inplace=True), nn.Linear(64, n_classes))
    if name == "wide_mlp_ae":
        layers = []
        for _ in range(3):
            layers += [nn.Linear(4096, 4096), nn.Tanh()]
        layers.append(nn.Linear(4096, 4096))
        return nn.Sequential(*layers)
    raise ValueError("no torch stand-in for model %r" % name)


class TorchArm(object):
    """One rank of the stock-torch training loop on a device-resident synthetic dataset."""

    def __init__(self, model, B, local_rank, world, n_train=8, lr=0.01, ddp=True):
        import torch
        import torch.nn as nn
        self.torch = torch
        self.name = model
        self.B = B
        self.dev = torch.device("cuda", local_rank)
        self.image = model in ("alexnet", "cifar_convnet")
        torch.backends.cudnn.benchmark = True
        net = build_torch_model(model).to(self.dev)
        if self.image:
            net = net.to(memory_format=torch.channels_last)
        self.net = net
        self.module = net
        if world > 1 and ddp:
            from torch.nn.parallel import DistributedDataParallel
            self.net = DistributedDataParallel(net, device_ids=[local_rank], gradient_as_bucket_view=True)
        self.opt = torch.optim.SGD(net.parameters(), lr=lr, momentum=0.9, weight_decay=0.0005)
        self.loss_fn = nn.CrossEntropyLoss() if self.image else nn.MSELoss()
        g = torch.Generator(device=self.dev).manual_seed(1234 + local_rank)
        shape = {"alexnet": (227, 227, 3), "cifar_convnet": (32, 32, 3), "wide_mlp_ae": (4096,)}[model]
        self.shape = shape
        n = n_train * B
        if self.image:
            self.data = torch.randint(0, 256, (n,) + shape, dtype=torch.uint8, device=self.dev, generator=g)
            self.labels = torch.randint(0, 1000 if model == "alexnet" else 10, (n,), device=self.dev, generator=g)
        else:
            self.data = torch.rand((n,) + shape, device=self.dev, generator=g) * 2 - 1
            self.labels = None
        self.perm = torch.randperm(n, device=self.dev, generator=g)
        self.pos = 0
        self.last_loss = None

    def batch_resident(self):
        torch = self.torch
        n = self.perm.numel()
        if self.pos + self.B > n:
            self.pos = 0
        idx = self.perm[self.pos:self.pos + self.B]
        self.pos += self.B
        return self.prepare(self.data[idx], self.labels[idx] if self.labels is not None else None)

    def prepare(self, x, y):
        torch = self.torch
        if self.image:
            x = (x.permute(0, 3, 1, 2).float() - 127.5) * (1.0 / 127.5)   # NHWC uint8 -> logical NCHW, channels_last memory
            x = x.contiguous(memory_format=torch.channels_last)
            return x, y.long()
        return x, x

    def train(self, x, y):
        torch = self.torch
        self.opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = self.net(x)
        loss = self.loss_fn(out.float(), y)
        loss.backward()
        self.opt.step()
        self.last_loss = loss.detach()
        return self.last_loss

    def step_resident(self):
        return self.train(*self.batch_resident())


def _timed_loop(args, world, fn, sync_all, max_over_ranks, min_seconds=0.0):
    """Time exactly args.steps calls of ``fn`` (more, in whole multiples, until min_seconds) with CUDA events."""
    import torch
    steps = args.steps
    total_ms, done = 0.0, 0
    while True:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        sync_all()
        total_ms += max_over_ranks(e0.elapsed_time(e1))
        done += steps
        if total_ms >= min_seconds * 1000.0:
            return total_ms, done


def run_torch(args, world, rank, local_rank, sync_all, max_over_ranks, ClockSampler):
    import numpy as np
    import torch
    B = args.batch
    arm = TorchArm(args.model, B, local_rank, world)
    warm = max(args.warmup, 3)
    for _ in range(warm):
        arm.step_resident()
    sync_all()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    ms_total, steps = _timed_loop(args, world, arm.step_resident, sync_all, max_over_ranks, args.min_seconds)
    clocks = sampler.stop() if sampler is not None else None
    ms = ms_total / steps
    loss_after = float(arm.last_loss.item())
    # end to end: pinned host uint8 minibatch -> device each step, loss read back each step
    e2e = None
    if not args.skip_e2e:
        rng = np.random.RandomState(100 + rank)
        pool = []
        for _ in range(4):
            if arm.image:
                d = torch.from_numpy(rng.randint(0, 256, size=(B,) + arm.shape).astype(np.uint8)).pin_memory()
                lab = torch.from_numpy(rng.randint(0, 1000 if args.model == "alexnet" else 10, size=B)
                                       .astype(np.int32)).pin_memory()
            else:
                d = torch.from_numpy(rng.uniform(-1, 1, size=(B,) + arm.shape).astype(np.float32)).pin_memory()
                lab = None
            pool.append((d, lab))
        copy_stream = torch.cuda.Stream()
        h2d = pool[0][0].numel() * pool[0][0].element_size() + (B * 4 if arm.image else 0)

        def upload(i):
            d, lab = pool[i % 4]
            with torch.cuda.stream(copy_stream):
                xd = d.to(arm.dev, non_blocking=True)
                ld = lab.to(arm.dev, non_blocking=True) if lab is not None else None
                ev = torch.cuda.Event()
                ev.record(copy_stream)
            return xd, ld, ev

        def e2e_steps(n):
            nxt = upload(0)
            last = None
            for i in range(n):
                xd, ld, ev = nxt
                nxt = upload(i + 1)            # double buffering, like Trainer.feed/step(prefetched=True)
                torch.cuda.current_stream().wait_event(ev)
                loss = arm.train(*arm.prepare(xd, ld))
                last = float(loss.item())
            return last

        e2e_steps(warm)
        sync_all()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = e2e_steps(args.steps)
        e1.record()
        sync_all()
        wall = (time.perf_counter() - t0) * 1000.0
        e2e_ms = max(max_over_ranks(e0.elapsed_time(e1)), max_over_ranks(wall)) / args.steps
        e2e = {"value": world * B * 1000.0 / e2e_ms, "unit": "images/s", "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4, "last_loss": last,
               "api": "torch.nn module + torch.optim.SGD, pinned host -> device each step, loss.item() each step"}
    if rank == 0:
        print(json.dumps({
            "metric": "AlexNet images/sec" if args.model == "alexnet" else "%s samples/sec" % args.model,
            "value": world * B * 1000.0 / ms, "unit": "images/s", "n_gpus": world, "steps": steps, "warmup": warm,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic", "impl": "torch", "stand_in": True,
            "config": {"model": args.model, "global_batch": world * B, "per_gpu_batch": B,
                       "parallelism": "dp%d" % world if world > 1 else "single",
                       "grad_exchange": "none" if world == 1 else "DistributedDataParallel (NCCL all-reduce) + SGD",
                       "cuda_graph": False,
                       "precision": "bf16 autocast, channels_last, fp32 master weights, cuDNN/cuBLAS kernels",
                       "l2_policy": "every step gathers a new minibatch from the resident dataset"},
            "gpu_launches": 0, "clocks": clocks, "e2e": e2e, "loss": loss_after,
            "note": "stand-in, not the reference: stock torch.nn / cuDNN / cuBLAS / NCCL; no veles_b200 code on this path"}))


# ------------------------------------------------------------------------------------------------------------------
# Reference-protocol emulation
# ------------------------------------------------------------------------------------------------------------------
class _Master(threading.Thread):
    """The reference's master: one Python thread answering job requests and applying updates under a lock."""

    def __init__(self, endpoint, state, n_slaves):
        super(_Master, self).__init__(daemon=True)
        import zmq
        self.ctx = zmq.Context.instance()
        self.sock = self.ctx.socket(zmq.ROUTER)
        self.sock.bind(endpoint)
        self.state = state            # name -> numpy array (the master's weights)
        self.lock = threading.Lock()
        self.n_slaves = n_slaves
        self.done = 0
        self.updates = 0
        self.bytes_in = 0
        self.bytes_out = 0

    def run(self):
        import zmq
        while self.done < self.n_slaves:
            ident, _, kind, payload = self.sock.recv_multipart()
            if kind == b"job":
                with self.lock:       # generate_data_for_slave: pickle the CURRENT weights into the job
                    blob = pickle.dumps(self.state, protocol=pickle.HIGHEST_PROTOCOL)
                self.bytes_out += len(blob)
                self.sock.send_multipart([ident, b"", blob])
            elif kind == b"update":
                delta = pickle.loads(payload)
                self.bytes_in += len(payload)
                with self.lock:       # apply_data_from_slave under the workflow lock
                    for k, v in delta.items():
                        self.state[k] += v
                    self.updates += 1
                self.sock.send_multipart([ident, b"", b"ok"])
            elif kind == b"bye":
                self.done += 1
                self.sock.send_multipart([ident, b"", b"ok"])
        self.sock.close(0)


def run_ps_emulation(args, world, rank, local_rank, sync_all, max_over_ranks, ClockSampler):
    import numpy as np
    import torch
    import zmq
    B = args.batch
    arm = TorchArm(args.model, B, local_rank, world, ddp=False)
    names = [n for n, _ in arm.module.named_parameters()]
    params = dict(arm.module.named_parameters())
    endpoint = "ipc:///tmp/veles_b200_ps_%s.ipc" % os.environ.get("MASTER_PORT", str(os.getppid()))
    master = None
    if rank == 0:
        state = {n: params[n].detach().float().cpu().numpy().copy() for n in names}
        master = _Master(endpoint, state, world)
        master.start()
    sync_all()
    ctx = zmq.Context.instance()

    def connect():
        s = ctx.socket(zmq.REQ)
        s.connect(endpoint)
        return s

    sock = connect()
    sock2 = connect() if args.ps_async else None
    host_bufs = {n: torch.empty_like(params[n], device="cpu").pin_memory() for n in names}
    job_weights = {}

    def fetch_job(s):
        s.send_multipart([b"job", b""])
        return pickle.loads(s.recv())

    def apply_job(job):
        # apply_data_from_master: host arrays -> device (Array.map_write / unmap)
        with torch.no_grad():
            for n in names:
                host_bufs[n].copy_(torch.from_numpy(job[n]))
                params[n].copy_(host_bufs[n], non_blocking=True)
        job_weights.clear()
        job_weights.update(job)

    # --ps-async: at most two jobs in flight - while job i trains, a helper thread already requests job i+1 (whose
    # weights are then one update stale, as with the reference's --async-slave)
    prefetch = {}

    def one_step():
        job = prefetch.pop("job", None)
        if job is None:
            job = fetch_job(sock)
        apply_job(job)
        th = None
        if args.ps_async:
            th = threading.Thread(target=lambda: prefetch.__setitem__("job", fetch_job(sock2)))
            th.start()
        arm.step_resident()
        delta = {}
        with torch.no_grad():                          # generate_data_for_master: map_read + delta against the job
            for n in names:
                delta[n] = params[n].detach().float().cpu().numpy() - job_weights[n]
        sock.send_multipart([b"update", pickle.dumps(delta, protocol=pickle.HIGHEST_PROTOCOL)])
        sock.recv()
        if th is not None:
            th.join()

    warm = max(args.warmup, 3)
    for _ in range(warm):
        one_step()
    sync_all()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one_step()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1000.0
    sync_all()
    clocks = sampler.stop() if sampler is not None else None
    ms = max_over_ranks(wall_ms) / args.steps
    loss_after = float(arm.last_loss.item())
    sock.send_multipart([b"bye", b""])
    sock.recv()
    sock.close(0)
    if sock2 is not None:
        sock2.close(0)
    if rank == 0:
        master.join(timeout=30)
        nbytes = sum(int(np.prod(p.shape)) * 4 for p in params.values())
        print(json.dumps({
            "metric": "AlexNet images/sec" if args.model == "alexnet" else "%s samples/sec" % args.model,
            "value": world * B * 1000.0 / ms, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic", "impl": "ps_emulation", "stand_in": True,
            "config": {"model": args.model, "global_batch": world * B, "per_gpu_batch": B,
                       "parallelism": "master + %d slave%s" % (world, "" if world == 1 else "s"),
                       "grad_exchange": "pickled weights down / pickled deltas up over ZeroMQ ipc, master applies under a lock",
                       "jobs_in_flight": 2 if args.ps_async else 1, "payload_bytes_each_way": nbytes,
                       "timing": "wall clock around the slave loop (the exchange is host work), max over ranks"},
            "gpu_launches": 0, "clocks": clocks, "e2e": None, "loss": loss_after,
            "master": {"updates": master.updates, "bytes_in": master.bytes_in, "bytes_out": master.bytes_out},
            "note": "reference-equivalent (emulated), not the reference's own number: the master-slave schedule of "
                    "veles/server.py + client.py + workflow.py with torch/cuDNN compute on the slaves"}))
