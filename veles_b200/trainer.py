"""Trainer: the imperative public API around a StandardWorkflow.

    trainer = veles_b200.Trainer("alexnet", minibatch_size=256)
    loss = trainer.step(images_uint8, labels_int32)       # host tensors in, python float out

``step`` uploads the minibatch from (pinned) host memory on a side stream, runs the whole
training iteration (forward, evaluator, backward, fused GD update) as a replayed CUDA graph and
reads the loss back — this is what ``bench.py`` times for its end-to-end number.
``step_resident`` trains on the device-resident synthetic dataset (kernel-only number).
"""
from __future__ import annotations

import os


class Trainer(object):
    def __init__(self, model="alexnet", minibatch_size=None, device=None, use_graph=True, streaming=False,
                 data_parallel=None, n_train=None, layers_kwargs=None, loader_extra=None, seed=1234):
        import torch
        from . import prng
        from .backends import Device
        from .config import root
        from .models.configs import MODEL_CONFIGS, build_workflow
        from .parallel.step import FlatStep
        root.common.disable.snapshotting = True
        prng.get().seed(seed)
        self.model = model
        self.cfg = MODEL_CONFIGS[model]
        self.dp = data_parallel
        self.rank = data_parallel.rank if data_parallel is not None else 0
        self.world = data_parallel.world_size if data_parallel is not None else 1
        self.minibatch_size = minibatch_size or self.cfg["minibatch"]
        self.device = device or Device()
        extra = dict(loader_extra or {})
        kwargs = {}
        if streaming:
            from .loader import stream  # noqa: F401
            extra.setdefault("dtype", "uint8" if len(self.cfg["sample_shape"]) == 3 else "float32")
            if extra["dtype"] == "uint8":
                import numpy
                n = int(numpy.prod(self.cfg["sample_shape"]))
                extra.setdefault("mean", numpy.full(n, 127.5, dtype=numpy.float32))
                extra.setdefault("rdisp", numpy.full(n, 1.0 / 127.5, dtype=numpy.float32))
        else:
            if len(self.cfg["sample_shape"]) == 3:
                extra.setdefault("normalization_type", "mean_disp")
        # data parallel: the loader serves GLOBAL minibatches of world*B and materialises this rank's B rows
        self.wf = build_workflow(model, minibatch_size=self.minibatch_size * self.world,
                                 n_train=(n_train or 8 * self.minibatch_size) * self.world, n_valid=0,
                                 max_epochs=None, loader_extra=extra, layers_kwargs=layers_kwargs,
                                 shard_rank=self.rank, shard_world=self.world, **kwargs) if not streaming else \
            self._build_streaming(model, extra, layers_kwargs)
        if data_parallel is not None and streaming:
            data_parallel.grad_scale = 1.0 / self.world  # every rank divided by its local batch size
        if data_parallel is not None:
            data_parallel.attach(self.wf)
        self.wf.initialize(device=self.device)
        if data_parallel is not None:
            data_parallel.finalize(self.wf)
        self.streaming = streaming
        self.stepper = FlatStep(self.wf, use_graph=use_graph and self.device.exists, comm=data_parallel)
        self._loss_host = None
        if self.device.exists:
            self._loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
        self.last_h2d_bytes = 0

    def _build_streaming(self, model, extra, layers_kwargs):
        from .dummy import DummyLauncher
        from .znicz.standard_workflow import StandardWorkflow
        cfg = self.cfg
        loader_cfg = {"minibatch_size": self.minibatch_size, "sample_shape": cfg["sample_shape"],
                      "n_classes": cfg["n_classes"] or 10}
        if cfg["loss_function"] == "mse":
            loader_cfg.setdefault("targets", "input")   # the MSE benchmark models are autoencoders
        loader_cfg.update(extra)
        return StandardWorkflow(DummyLauncher(), layers=cfg["layers"](**(layers_kwargs or {})),
                                loader_name="stream", loader_config=loader_cfg,
                                loss_function=cfg["loss_function"], decision_config={"max_epochs": None})

    # -- steps --------------------------------------------------------------------------------------------------------
    def step_resident(self):
        """One training iteration on the device-resident dataset; nothing is synchronised."""
        self.stepper.run(train=True, serve=True)

    def feed(self, data, labels=None, targets=None):
        self.last_h2d_bytes = self.wf.loader.feed(data, labels, targets) if targets is not None else \
            self.wf.loader.feed(data, labels)
        return self.last_h2d_bytes

    def step(self, data=None, labels=None, prefetched=False):
        """Train on one host minibatch and return its mean loss (forces one D2H read)."""
        import torch
        if not prefetched:
            self.feed(data, labels)
        self.stepper.run(train=True, serve=True)
        if os.environ.get("VELES_HEARTBEAT_FILE"):
            from .parallel.supervisor import heartbeat
            heartbeat()
        return self.read_loss()

    def step_pipelined(self, data=None, labels=None, prefetched=False):
        """Like :meth:`step`, but the host does not wait for THIS step: the loss is copied to pinned memory
        asynchronously and the call returns the mean loss of the PREVIOUS pipelined step (None the first time), whose
        copy has long finished.  The host therefore enqueues step i+1 while the device still runs step i - the
        ~0.15 ms of per-step host work (loader bookkeeping, graph launch) no longer sits between two steps.  Every
        step's loss still crosses to the host; :meth:`flush_loss` returns the last one."""
        if not prefetched:
            self.feed(data, labels)
        self.stepper.run(train=True, serve=True)
        if os.environ.get("VELES_HEARTBEAT_FILE"):
            from .parallel.supervisor import heartbeat
            heartbeat()
        prev = self._take_pending_loss()
        self._post_loss_read()
        return prev

    def flush_loss(self):
        """Mean loss of the last :meth:`step_pipelined` call (waits for it), or None when nothing is pending."""
        return self._take_pending_loss()

    def _post_loss_read(self):
        if not self.device.exists:
            self._pending_loss = ("host", self.read_loss())
            return
        import torch
        ev = self.wf.evaluator
        src = ev.loss_sum.devmem if hasattr(ev, "loss_sum") else ev.metrics.devmem[:1]
        if self.__dict__.get("_loss_ring") is None:
            self._loss_ring = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
            self._loss_events = [torch.cuda.Event() for _ in range(2)]
            self._loss_slot = 0
        k = self._loss_slot = self._loss_slot ^ 1
        self._loss_ring[k].copy_(src, non_blocking=True)
        src.zero_()
        self._loss_events[k].record()
        self._pending_loss = ("device", k, self._loss_denominator())

    def _take_pending_loss(self):
        pending = self.__dict__.get("_pending_loss")
        self._pending_loss = None
        if pending is None:
            return None
        if pending[0] == "host":
            return pending[1]
        _, k, den = pending
        self._loss_events[k].synchronize()
        return float(self._loss_ring[k][0]) / den

    def _loss_denominator(self):
        """Samples behind the evaluator's loss sum on THIS rank (the resident loader serves global minibatches and
        materialises this rank's share)."""
        loader = self.wf.loader
        n = max(1, loader.minibatch_size)
        if not self.streaming and self.world > 1:
            n = max(1, int(getattr(loader, "local_minibatch_size", 0)) or n // self.world)
        return n

    def read_loss(self):
        import torch
        ev = self.wf.evaluator
        if not self.device.exists:
            src = ev.loss_sum if hasattr(ev, "loss_sum") else ev.metrics
            val = float(src.mem[0])
            src.mem[0] = 0
            return val / self._loss_denominator()
        src = ev.loss_sum.devmem if hasattr(ev, "loss_sum") else ev.metrics.devmem[:1]
        self._loss_host.copy_(src, non_blocking=True)
        src.zero_()
        torch.cuda.current_stream().synchronize()
        return float(self._loss_host[0]) / self._loss_denominator()

    def close(self):
        self.wf.launcher.stop()
