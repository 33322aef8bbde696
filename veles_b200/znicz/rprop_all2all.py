"""RPROP for fully connected layers (``rprop_all2all``): per-weight step sizes adapted by the sign agreement of
consecutive gradients (Riedmiller & Braun); forward is a plain All2All."""
from __future__ import annotations

import numpy

from ..memory import Array
from .all2all import All2All
from .gd import GradientDescent


class RPropAll2All(All2All):
    """Fully connected layer trained by resilient propagation."""
    TRANSPOSED_OK = False
    MAPPING = {"rprop_all2all"}


class GDRProp(GradientDescent):
    """RProp update: per-weight step sizes grown / shrunk by the sign history of the gradient."""
    MAPPING = {"rprop_all2all"}

    def __init__(self, workflow, **kwargs):
        super(GDRProp, self).__init__(workflow, **kwargs)
        self.initial_delta = kwargs.get("initial_delta", 0.01)
        self.min_delta, self.max_delta = kwargs.get("min_delta", 1e-6), kwargs.get("max_delta", 1.0)
        self.increase, self.decrease = kwargs.get("increase", 1.2), kwargs.get("decrease", 0.5)
        self.deltas = [Array(), Array()]
        self.prev_grads = [Array(), Array()]

    def numpy_update(self, w, g, vel, acc, lr, wd, l12, moment):
        slot = 0 if w.ndim == 2 else 1
        d, pg = self.deltas[slot], self.prev_grads[slot]
        if not d or d.shape != w.shape:
            d.reset(numpy.full(w.shape, self.initial_delta, dtype=numpy.float32))
            pg.reset(numpy.zeros(w.shape, dtype=numpy.float32))
        sign = numpy.sign(g * pg.mem)
        d.mem[...] = numpy.clip(numpy.where(sign > 0, d.mem * self.increase,
                                            numpy.where(sign < 0, d.mem * self.decrease, d.mem)),
                                self.min_delta, self.max_delta)
        g_eff = numpy.where(sign < 0, 0, g)
        w -= numpy.sign(g_eff) * d.mem
        pg.mem[...] = g_eff

    def device_update(self, forward=None):
        """RPROP state lives on the host (it is a small-model, full-batch method): update there, then upload."""
        self.gradient_weights.map_read()
        self.weights.map_write()
        self.numpy_update(self.weights.mem, self.gradient_weights.mem, None, None, 0, 0, 0, 0)
        self.weights.unmap()
        self.gradient_weights.map_write()
        self.gradient_weights.mem[...] = 0
        self.gradient_weights.unmap()
        if self.bias is not None and self.bias:
            self.gradient_bias.map_read()
            self.bias.map_write()
            self.numpy_update(self.bias.mem, self.gradient_bias.mem, None, None, 0, 0, 0, 0)
            self.bias.unmap()
            self.gradient_bias.map_write()
            self.gradient_bias.mem[...] = 0
            self.gradient_bias.unmap()
        if self.forward is not None and self.forward.weights_bf16_ is not None:
            self.forward.refresh_shadow()
