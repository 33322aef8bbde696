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


class GraphedChain(object):
    """A run of device units of a workflow that is walked by the ordinary unit scheduler (``Workflow.run()``, the
    ``veles`` command line) but executes as ONE replayed CUDA graph.

    The scheduler still visits every unit, evaluates its gates and notifies its successors; only the ``run()`` calls
    change: the chain's first unit runs (or replays) the whole chain, the others find their work already done.  A
    chain is valid for one signature - minibatch size, the GD hyper-parameters, the data-parallel gradient parity -
    and falls back to running its units one by one whenever capture fails (a unit that synchronises or allocates).
    This is what replaces the per-minibatch cost of the reference's scheduler on the public CLI path
    (``/root/reference/veles/units.py:485-505, 782-803``): gates keep deciding WHAT runs at class / epoch boundaries,
    the steady-state sequence inside a class is resolved once and replayed.
    """

    WARMUP = 2

    def __init__(self, workflow, units, name="chain"):
        self.wf = workflow
        self.units = list(units)
        self.name = name
        self.graphs = {}
        self.eager_runs = {}
        self.enabled = True
        self.replays = 0
        self.captures = 0
        self._done = set()
        self.sync = {}         # key -> {"unmaps": [Array], "marks": [Array]} learnt during the eager / capture runs
        for u in self.units:
            u.graph_chain_ = self

    def _signature(self):
        wf = self.wf
        hyper = tuple((u.learning_rate, u.learning_rate_bias, u.weights_decay, u.gradient_moment,
                       getattr(u, "weights_decay_bias", 0.0), getattr(u, "gradient_moment_bias", 0.0))
                      for u in self.units if hasattr(u, "learning_rate"))
        dp = getattr(wf, "dp_group_", None)
        loader = getattr(wf, "loader", None)
        return (int(loader.minibatch_size) if loader is not None else 0,
                int(getattr(loader, "minibatch_class", 0) or 0) if loader is not None else 0,
                hyper, getattr(dp, "parity", 0))

    def _run_all(self, key=None):
        """Run the members one by one, logging the Arrays they upload (``unmap`` of host-written data) and the ones
        their kernels write (``mark_device_written``): a replay repeats the kernels only, so the chain uploads the
        former before and flags the latter after every replay."""
        from .. import memory
        if key is None:
            for u in self.units:
                u.run()
            return
        log = {"unmaps": [], "marks": []}
        memory.SYNC_TRACE.log = log
        try:
            for u in self.units:
                u.run()
        finally:
            memory.SYNC_TRACE.log = None
        rec = self.sync.setdefault(key, {"unmaps": [], "marks": []})
        for kind in ("unmaps", "marks"):
            seen = set(id(a) for a in rec[kind])
            for a in log[kind]:
                if id(a) not in seen:
                    seen.add(id(a))
                    rec[kind].append(a)

    def _replay(self, key, g):
        rec = self.sync.get(key)
        if rec is not None:
            for a in rec["unmaps"]:
                a.unmap()
        g.replay()
        if rec is not None:
            for a in rec["marks"]:
                a.mark_device_written()
        self.replays += 1

    def run_member(self, unit):
        """Called by the scheduler instead of ``unit.run()``."""
        if unit is not self.units[0]:
            if id(unit) in self._done:
                self._done.discard(id(unit))      # executed with the head of the chain
                return
            unit.run()                            # the head was gated off this pass: plain behaviour
            return
        if not self.enabled:
            unit.run()
            return
        import torch
        self._done = set(id(u) for u in self.units[1:] if not u.gate_skip and not u.gate_block)
        # every member must be due this pass; otherwise run unit by unit
        if len(self._done) != len(self.units) - 1:
            self._done = set()
            unit.run()
            return
        key = self._signature()
        g = self.graphs.get(key)
        if g is not None:
            self._replay(key, g)
            return
        n = self.eager_runs.get(key, 0)
        if n < self.WARMUP:
            self.eager_runs[key] = n + 1
            self._run_all(key)
            return
        try:
            for a in self.sync.get(key, {}).get("unmaps", ()):
                a.unmap()          # uploads happen OUTSIDE the captured region
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._run_all(key)
        except Exception as e:   # a unit that cannot be captured: stay correct, stay eager
            self.enabled = False
            self.wf.warning("CUDA-graph replay of the %s chain is off (%s: %s)", self.name, type(e).__name__, e)
            torch.cuda.synchronize()
            self._run_all()
            return
        self.captures += 1
        self.graphs[key] = g
        if len(self.graphs) > 8:
            self.graphs.pop(next(iter(self.graphs)))
        self._replay(key, g)
        if self.captures >= 16 and self.replays < 4 * self.captures:
            # the signature changes nearly every minibatch (a per-iteration learning-rate schedule): capturing
            # costs more than it saves
            self.enabled = False
            self.graphs.clear()

    def invalidate(self):
        self.graphs.clear()
        self.eager_runs.clear()


def install_graph_replay(wf):
    """Chains of a StandardWorkflow on a CUDA device: [forwards..., evaluator] and [GD units in backward order].
    Returns the chains (empty when the workflow does not qualify).  ``root.common.engine.graph_replay = False``
    switches it off."""
    from ..config import root, get
    device = getattr(wf, "device", None)
    if device is None or not getattr(device, "exists", False) or not get(root.common.engine.graph_replay, True):
        return []
    if getattr(wf, "is_slave", False) or getattr(wf, "is_master", False):
        return []      # job-level master-slave runs exchange host arrays around every minibatch
    forwards, gds = list(getattr(wf, "forwards", []) or []), list(getattr(wf, "gds", []) or [])
    evaluator = getattr(wf, "evaluator", None)
    chains = []

    def consecutive(units):
        """Each unit must be the only successor-in-chain of its predecessor: the scheduler then visits them in order."""
        for a, b in zip(units, units[1:]):
            if a not in b.links_from:
                return False
        return all(getattr(u, "on_device", True) is not False for u in units)

    fwd_units = forwards + ([evaluator] if evaluator is not None else [])
    if len(fwd_units) > 1 and consecutive(fwd_units):
        chains.append(GraphedChain(wf, fwd_units, "forward"))
    bwd_units = [g for g in reversed(gds) if g is not None]
    if len(bwd_units) > 1 and consecutive(bwd_units):
        chains.append(GraphedChain(wf, bwd_units, "backward"))
    wf.graph_chains_ = chains
    return chains
