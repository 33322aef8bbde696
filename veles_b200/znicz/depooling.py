"""Depooling (``depooling`` layer type): scatters a pooled tensor back to the positions recorded by a max-like
pooling unit (``output_offset`` = that unit's ``input_offset``); zeros elsewhere.  Backward gathers."""
from __future__ import annotations

import numpy

from .nn_units import ForwardBase, GDBase


class Depooling(ForwardBase):
    """Puts pooled values back at the positions a max pooling layer took them from."""
    hide_from_registry = False
    MAPPING = {"depooling"}

    def __init__(self, workflow, **kwargs):
        super(Depooling, self).__init__(workflow, **kwargs)
        self.output_offset = None         # link to Pooling.input_offset
        self.output_shape_source = kwargs.get("output_shape_source")   # link to Pooling.input
        self.demand("output_offset", "output_shape_source")

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input or not self.output_offset or not self.output_shape_source:
            return True
        self.device = device
        self.alloc_output(self.output_shape_source.shape)
        if self.on_device:
            self.output.dev_dtype = self.input.devmem.dtype
            self.output.initialize(self.device_)
        self._backend_init_()
        return None

    def numpy_run(self):
        self.input.map_read()
        self.output_offset.map_read()
        self.output.map_invalidate()
        self.output.mem[...] = 0
        self.output.mem.reshape(-1)[self.output_offset.mem.reshape(-1)] = self.input.mem.reshape(-1)

    def b200_run(self):
        out = self.output.devmem
        out.zero_()
        out.view(-1)[self.output_offset.devmem.view(-1).long()] = self.input.devmem.reshape(-1).to(out.dtype)
        self.output.mark_device_written()


class GDDepooling(GDBase):
    """Backward pass of Depooling."""
    hide_from_registry = False
    MAPPING = {"depooling"}

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
        f.output_offset.map_read()
        self.err_input.map_invalidate()
        self.err_input.mem[...] = self.err_output.mem.reshape(-1)[f.output_offset.mem.reshape(-1)].reshape(
            self.err_input.shape)

    def b200_run(self):
        f = self.forward
        ei = self.err_input.devmem
        ei.view(-1).copy_(self.err_output.devmem.reshape(-1)[f.output_offset.devmem.view(-1).long()])
        self.err_input.mark_device_written()
