"""Flat step scheduler: the steady-state training iteration of a StandardWorkflow as one
pre-resolved unit list, optionally captured into a CUDA graph.

The reference walks the unit graph through 3-5 Python wrappers, two locks and a Bool
evaluation per unit per minibatch (``veles/units.py:485-505, 782-885``) — fine at
2014 GPU speeds, dominant at B200 speeds.  Here the gate logic still decides WHAT runs (the
Decision/Loader Bools are evaluated on class/epoch boundaries), but inside a class the same
chain of device units runs every minibatch, so it is resolved once and replayed:

    loader gather (eager: its offset changes)  ->  [ forwards -> evaluator -> GDs ]  (CUDA graph)

The graph is invalidated whenever anything baked into it changes (learning rates from an LR
schedule, minibatch class, batch size of the tail minibatch).
"""
from __future__ import annotations


class FlatStep(object):
    def __init__(self, workflow, use_graph=True, comm=None):
        self.wf = workflow
        self.use_graph = use_graph
        self.comm = comm              # DataParallel group or None
        self.graphs = {}              # key -> torch.cuda.CUDAGraph
        self.eager_steps = {}         # key -> number of eager runs so far (warm-up before capture)
        self.warmup = 2

    # -- what runs ---------------------------------------------------------------------------------------------
    def _signature(self, train):
        wf = self.wf
        hyper = tuple((gd.learning_rate, gd.learning_rate_bias, gd.weights_decay, gd.gradient_moment)
                      for gd in wf.gds if hasattr(gd, "learning_rate"))
        return (bool(train), int(wf.loader.minibatch_size), hyper, getattr(self.comm, "parity", 0))

    def _body(self, train):
        wf = self.wf
        for f in wf.forwards:
            f.run()
        wf.evaluator.run()
        if train:
            for gd in reversed(wf.gds):
                gd.run()
            if self.comm is not None:
                self.comm.finish_step()

    def run(self, train=True, serve=True):
        """One minibatch: serve it (eager) then run/replay the compute chain."""
        import torch
        wf = self.wf
        if serve:
            wf.loader.run()
        if not self.use_graph:
            self._body(train)
            return
        key = self._signature(train)
        g = self.graphs.get(key)
        if g is not None:
            g.replay()
            return
        n = self.eager_steps.get(key, 0)
        if n < self.warmup:
            self.eager_steps[key] = n + 1
            self._body(train)
            return
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._body(train)
        self.graphs[key] = g
        if len(self.graphs) > 8:
            self.graphs.pop(next(iter(self.graphs)))
        g.replay()

    def invalidate(self):
        self.graphs.clear()
        self.eager_steps.clear()
