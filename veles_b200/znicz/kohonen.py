"""Kohonen self-organising map: ``KohonenForward`` (winner per sample) and ``KohonenTrainer`` (neighbourhood
update) — listed among the Znicz algorithms (``docs/source/manualrst_veles_algorithms.rst:71-112``).  Host/torch
implementation: the map is small and this path is outside the benchmarked hot loop."""
from __future__ import annotations

import numpy

from ..accelerated_units import AcceleratedUnit
from ..distributable import TriviallyDistributable
from ..memory import Array
from .. import prng


class KohonenBase(AcceleratedUnit, TriviallyDistributable):
    hide_from_registry = True

    def numpy_init(self):
        pass

    def b200_init(self):
        pass


class KohonenForward(KohonenBase):
    """Winner lookup of a Kohonen self-organising map."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(KohonenForward, self).__init__(workflow, **kwargs)
        self.output = Array(shallow_pickle=True)
        self.weights = None
        self.demand("input", "weights")

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input or self.weights is None or not self.weights:
            return True
        self.device = device
        self.output.reset(numpy.zeros(self.input.shape[0], dtype=numpy.int32))
        return None

    def numpy_run(self):
        self.input.map_read()
        self.weights.map_read()
        x = self.input.mem.reshape(self.input.shape[0], -1)
        w = self.weights.mem.reshape(self.weights.shape[0], -1)
        d = (x * x).sum(1)[:, None] - 2 * x.dot(w.T) + (w * w).sum(1)[None, :]
        self.output.map_invalidate()
        self.output.mem[...] = d.argmin(axis=1)

    b200_run = numpy_run


class KohonenTrainer(KohonenBase):
    """Kohonen map training: neighbourhood-weighted pull of the weights toward each sample."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(KohonenTrainer, self).__init__(workflow, **kwargs)
        self.shape = tuple(kwargs.get("shape", (8, 8)))
        self.gradient_decay = kwargs.get("gradient_decay", lambda t: 0.5 / (1.0 + t * 0.02))
        self.radius_decay = kwargs.get("radius_decay", lambda t: 1.0 / (1.0 + t * 0.2))
        self.weights = Array()
        self.winners = Array()
        self.time = 0
        self.rand = kwargs.get("rand", prng.get())
        self.demand("input")

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        n = self.shape[0] * self.shape[1]
        if not self.weights:
            w = numpy.zeros((n, self.input.sample_size), dtype=numpy.float32)
            self.rand.fill(w, -0.05, 0.05)
            self.weights.reset(w)
        self.winners.reset(numpy.zeros(n, dtype=numpy.int32))
        ys, xs = numpy.divmod(numpy.arange(n), self.shape[1])
        self.coords_ = numpy.stack([ys, xs], axis=1).astype(numpy.float32)
        return None

    def numpy_run(self):
        self.input.map_read()
        self.weights.map_write()
        x = self.input.mem.reshape(self.input.shape[0], -1)
        w = self.weights.mem
        lr = self.gradient_decay(self.time)
        radius = max(self.radius_decay(self.time) * max(self.shape) / 2.0, 0.5)
        d = (x * x).sum(1)[:, None] - 2 * x.dot(w.T) + (w * w).sum(1)[None, :]
        win = d.argmin(axis=1)
        numpy.add.at(self.winners.mem, win, 1)
        grid_d2 = ((self.coords_[win][:, None, :] - self.coords_[None, :, :]) ** 2).sum(-1)
        h = numpy.exp(-grid_d2 / (2 * radius * radius))            # [batch, neurons]
        hs = h.sum(0)[:, None]                                     # batch-SOM step towards the weighted mean
        w += lr * numpy.where(hs > 1e-12, h.T.dot(x) / numpy.maximum(hs, 1e-12) - w, 0.0)
        self.time += 1

    b200_run = numpy_run
