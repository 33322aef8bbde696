"""Dropout (``dropout`` layer type, kwarg ``dropout_ratio``).

Training minibatches: ``output = input * mask`` with ``mask ∈ {0, 1/(1-ratio)}``; validation /
test minibatches and ``forward_mode``: identity.  Backward multiplies by the same mask.
Device path: the mask is never stored — forward draws it from per-thread
xorshift128+ states inside the kernel, backward regenerates it from the saved pre-forward states.
"""
from __future__ import annotations

import numpy

from .. import prng
from ..loader.base import TRAIN
from ..memory import Array
from .nn_units import ForwardBase, GDBase


class DropoutForward(ForwardBase):
    """Zeroes a `dropout_ratio` share of the inputs (mask from in-kernel xorshift128+ states) and rescales the
    rest; identity in evaluation."""
    hide_from_registry = False
    MAPPING = {"dropout"}
    LOADER_ATTRS = ("minibatch_class",)
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0022"
    N_STATES = 16384

    def __init__(self, workflow, **kwargs):
        super(DropoutForward, self).__init__(workflow, **kwargs)
        self.dropout_ratio = float(kwargs.get("dropout_ratio", 0.5))
        if not 0.0 <= self.dropout_ratio < 1.0:
            raise ValueError("dropout_ratio must be in [0, 1)")
        self.rand = kwargs.get("rand", prng.get())
        self.mask = Array(shallow_pickle=True)
        self.states = Array()
        self.minibatch_class = TRAIN

    def init_unpickled(self):
        super(DropoutForward, self).init_unpickled()
        self.states_prev_ = None
        self.states_next_ = None
        self.active_ = True

    def package_export(self):
        return {"dropout_ratio": self.dropout_ratio}   # identity at inference time

    @property
    def training(self):
        return (not self.forward_mode) and int(self.minibatch_class) == TRAIN and not self.testing

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        self.alloc_output(self.input.shape)
        if not self.states:
            self.states.reset(self.rand.randint(1, 2 ** 62, size=(self.N_STATES, 2)).astype(numpy.int64))
        self._backend_init_()
        return None

    def numpy_init(self):
        if not self.mask or self.mask.shape != self.input.shape:
            self.mask.reset(numpy.ones(self.input.shape, dtype=numpy.float32))

    def numpy_run(self):
        self.input.map_read()
        self.output.map_invalidate()
        x = self.input.mem.astype(numpy.float32)
        self.active_ = self.training
        if not self.active_:
            self.output.mem[...] = x
            return
        keep = self.rand.random_sample(x.shape) >= self.dropout_ratio
        self.mask.mem[...] = keep.astype(numpy.float32) / (1.0 - self.dropout_ratio)
        self.output.mem[...] = x * self.mask.mem

    def b200_init(self):
        import torch
        dev = self.device_
        self.output.dev_dtype = self.input.devmem.dtype
        self.output.initialize(dev)
        self.states.initialize(dev)
        self.states_prev_ = torch.empty_like(self.states.devmem)

    def b200_run(self):
        from ..ops import kernels
        x = self.input.devmem
        y = self.output.devmem
        self.active_ = self.training
        if not self.active_:
            y.copy_(x.view(y.shape))
        else:
            xs = x if x.is_contiguous() else x.contiguous()
            self.states_prev_.copy_(self.states.devmem)
            if y.is_contiguous():
                kernels.dropout(xs, y, self.states_prev_, self.states.devmem, self.dropout_ratio)
            else:
                tmp = xs.clone()
                kernels.dropout(xs, tmp, self.states_prev_, self.states.devmem, self.dropout_ratio)
                y.copy_(tmp.view(y.shape))
            self.states.mark_device_written()
        self.output.mark_device_written()


class DropoutBackward(GDBase):
    """Applies the forward mask (regenerated from the saved generator states) to the gradient."""
    hide_from_registry = False
    MAPPING = {"dropout"}

    def initialize(self, device=None, **kwargs):
        if self.err_output is None or not self.err_output or self.input is None or not self.input:
            return True
        self.device = device
        self.alloc_err_input()
        if self.on_device:
            self.err_input.dev_dtype = self.device_.compute_dtype
            self.err_input.initialize(self.device_)
        self._backend_init_()
        return None

    def numpy_run(self):
        f = self.forward
        self.err_output.map_read()
        eo = self.err_output.mem.reshape(self.input.shape).astype(numpy.float32)
        self.err_input.map_invalidate()
        self.err_input.mem[...] = eo * f.mask.mem if f.active_ else eo

    def b200_run(self):
        from ..ops import kernels
        f = self.forward
        eo = self.err_output.devmem
        ei = self.err_input.devmem
        src = eo if (eo.is_contiguous() and eo.dtype == ei.dtype) else eo.contiguous().to(ei.dtype)
        if not f.active_:
            ei.copy_(src.view(ei.shape))
        elif ei.is_contiguous():
            kernels.dropout(src, ei, f.states_prev_, None, f.dropout_ratio)
        else:
            tmp = src.clone()
            kernels.dropout(src, tmp, f.states_prev_, None, f.dropout_ratio)
            ei.copy_(tmp.view(ei.shape))
        self.err_input.mark_device_written()
