"""Cutter: crops a rectangle out of NHWC images (``cutter`` layer type; kwargs ``padding`` = (left, top, right,
bottom) to cut away); backward pads the error with zeros.  Also ``Cutter1D`` for flat vectors."""
from __future__ import annotations

import numpy

from .nn_units import ForwardBase, GDBase


class Cutter(ForwardBase):
    """Crops a window (`padding` = left, top, right, bottom) out of an NHWC tensor."""
    hide_from_registry = False
    MAPPING = {"cutter"}

    def __init__(self, workflow, **kwargs):
        super(Cutter, self).__init__(workflow, **kwargs)
        self.padding = tuple(kwargs.get("padding", (0, 0, 0, 0)))
        if len(self.padding) != 4 or min(self.padding) < 0:
            raise ValueError("padding must be 4 non-negative numbers (left, top, right, bottom)")

    @property
    def in_shape4(self):
        s = self.input.shape
        return s if len(s) == 4 else s + (1,)

    def slices(self):
        B, H, W, C = self.in_shape4
        l, t, r, b = self.padding
        return slice(t, H - b), slice(l, W - r)

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        B, H, W, C = self.in_shape4
        l, t, r, b = self.padding
        if H - t - b < 1 or W - l - r < 1:
            raise ValueError("Nothing is left after cutting %s from %dx%d" % (self.padding, H, W))
        self.alloc_output((B, H - t - b, W - l - r, C))
        if self.on_device:
            self.output.dev_dtype = self.input.devmem.dtype
            self.output.initialize(self.device_)
        self._backend_init_()
        return None

    def numpy_run(self):
        self.input.map_read()
        sh, sw = self.slices()
        self.output.map_invalidate()
        self.output.mem[...] = self.input.mem.reshape(self.in_shape4)[:, sh, sw, :]

    def b200_run(self):
        sh, sw = self.slices()
        self.output.devmem.copy_(self.input.devmem.reshape(self.in_shape4)[:, sh, sw, :])
        self.output.mark_device_written()


class GDCutter(GDBase):
    """Backward pass of Cutter (zero-pads the gradient)."""
    hide_from_registry = False
    MAPPING = {"cutter"}

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
        sh, sw = f.slices()
        self.err_input.map_invalidate()
        ei = self.err_input.mem.reshape(f.in_shape4)
        ei[...] = 0
        ei[:, sh, sw, :] = self.err_output.mem.reshape(f.output.shape)

    def b200_run(self):
        f = self.forward
        sh, sw = f.slices()
        ei = self.err_input.devmem.reshape(f.in_shape4)
        ei.zero_()
        ei[:, sh, sw, :] = self.err_output.devmem.reshape(f.output.shape).to(ei.dtype)
        self.err_input.mark_device_written()


class Cutter1D(ForwardBase):
    """output[:, output_offset:output_offset+length] (*alpha) = input[:, input_offset:input_offset+length]"""
    hide_from_registry = False
    MAPPING = {"cutter_1d"}

    def __init__(self, workflow, **kwargs):
        super(Cutter1D, self).__init__(workflow, **kwargs)
        self.alpha = float(kwargs.get("alpha", 1.0))
        self.input_offset = int(kwargs.get("input_offset", 0))
        self.length = kwargs.get("length")

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        n = self.input.sample_size
        if self.length is None:
            self.length = n - self.input_offset
        if self.input_offset + self.length > n:
            raise ValueError("Cutter1D range exceeds the input sample size")
        self.alloc_output((self.input.shape[0], self.length))
        if self.on_device:
            self.output.dev_dtype = self.input.devmem.dtype
            self.output.initialize(self.device_)
        self._backend_init_()
        return None

    def numpy_run(self):
        self.input.map_read()
        x = self.input.mem.reshape(self.input.shape[0], -1)
        self.output.map_invalidate()
        self.output.mem[...] = self.alpha * x[:, self.input_offset:self.input_offset + self.length]

    def b200_run(self):
        x = self.input.devmem.reshape(self.input.shape[0], -1)
        self.output.devmem.copy_(x[:, self.input_offset:self.input_offset + self.length] * self.alpha)
        self.output.mark_device_written()


class GDCutter1D(GDBase):
    """Backward pass of Cutter1D."""
    hide_from_registry = False
    MAPPING = {"cutter_1d"}

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
        self.err_input.map_invalidate()
        ei = self.err_input.mem.reshape(self.input.shape[0], -1)
        ei[...] = 0
        ei[:, f.input_offset:f.input_offset + f.length] = f.alpha * self.err_output.mem.reshape(ei.shape[0], -1)

    def b200_run(self):
        f = self.forward
        ei = self.err_input.devmem.reshape(self.input.shape[0], -1)
        ei.zero_()
        ei[:, f.input_offset:f.input_offset + f.length] = \
            (self.err_output.devmem.reshape(ei.shape[0], -1) * f.alpha).to(ei.dtype)
        self.err_input.mark_device_written()
