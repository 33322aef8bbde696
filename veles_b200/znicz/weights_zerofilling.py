"""ZeroFiller (``zero_filter`` layer type): keeps a block-diagonal mask on a layer's weights so that a dense
conv/all2all layer behaves like a grouped one (kwarg ``grouping``); it re-applies the mask after every update."""
from __future__ import annotations

import numpy

from ..memory import Array
from .nn_units import ForwardBase, GDBase


class ZeroFiller(ForwardBase):
    """Keeps the cross-group weights of a grouped layer at zero after every update."""
    hide_from_registry = False
    MAPPING = {"zero_filter"}

    def __init__(self, workflow, **kwargs):
        super(ZeroFiller, self).__init__(workflow, **kwargs)
        self.grouping = int(kwargs.get("grouping", 1))
        self.mask = Array()
        self.weights = None
        self.target = kwargs.get("target")   # the Forward unit whose weights are masked

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        self.output = self.input   # pass-through
        tgt = self.target
        if tgt is not None and tgt.weights:
            self.weights = tgt.weights
            n_out, n_in = self.weights.shape[0], self.weights.size // self.weights.shape[0]
            if n_out % self.grouping or n_in % self.grouping:
                raise ValueError("weights %dx%d are not divisible by grouping %d" % (n_out, n_in, self.grouping))
            mask = numpy.zeros((n_out, n_in), dtype=numpy.float32)
            ro, ci = n_out // self.grouping, n_in // self.grouping
            for g in range(self.grouping):
                mask[g * ro:(g + 1) * ro, g * ci:(g + 1) * ci] = 1
            self.mask.reset(mask.reshape(self.weights.shape))
            if self.on_device:
                self.mask.initialize(self.device_)
        self._backend_init_()
        return None

    def numpy_run(self):
        if self.weights is None:
            return
        self.weights.map_write()
        self.weights.mem *= self.mask.mem

    def b200_run(self):
        if self.weights is None:
            return
        self.weights.devmem.mul_(self.mask.devmem.view(self.weights.devmem.shape))
        self.weights.mark_device_written()
        if self.target is not None and getattr(self.target, "weights_bf16_", None) is not None:
            self.target.refresh_shadow()


class GDZeroFiller(GDBase):
    """Backward pass-through."""
    hide_from_registry = False
    MAPPING = {"zero_filter"}

    def initialize(self, device=None, **kwargs):
        if self.err_output is None or not self.err_output:
            return True
        self.device = device
        self.err_input = self.err_output
        self._backend_init_()
        return None

    def numpy_run(self):
        pass

    def b200_run(self):
        pass
