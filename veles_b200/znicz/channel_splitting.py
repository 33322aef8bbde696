"""ChannelSplitter / ChannelMerger: NHWC <-> planar regrouping of channels (``channel_splitter``,
``channel_merger`` layer types): splitter turns [B,H,W,C] into [B,C,H,W] (each channel a separate plane),
merger is the inverse; their backward units apply the opposite permutation to the error."""
from __future__ import annotations

import numpy

from .nn_units import ForwardBase, GDBase


class _Permute(ForwardBase):
    hide_from_registry = True
    PERM = (0, 1, 2, 3)

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        if len(self.input.shape) != 4:
            raise ValueError("%s needs a 4-D input" % type(self).__name__)
        self.device = device
        self.alloc_output(tuple(self.input.shape[i] for i in self.PERM))
        if self.on_device:
            self.output.dev_dtype = self.input.devmem.dtype
            self.output.initialize(self.device_)
        self._backend_init_()
        return None

    def numpy_run(self):
        self.input.map_read()
        self.output.map_invalidate()
        self.output.mem[...] = self.input.mem.transpose(self.PERM)

    def b200_run(self):
        self.output.devmem.copy_(self.input.devmem.reshape(self.input.shape).permute(*self.PERM))
        self.output.mark_device_written()


class ChannelSplitter(_Permute):
    """Splits the channel axis of an NHWC tensor into groups (one output per group)."""
    hide_from_registry = False
    MAPPING = {"channel_splitter"}
    PERM = (0, 3, 1, 2)


class ChannelMerger(_Permute):
    """Concatenates tensors along the channel axis."""
    hide_from_registry = False
    MAPPING = {"channel_merger"}
    PERM = (0, 2, 3, 1)


class _GDPermute(GDBase):
    hide_from_registry = True
    PERM = (0, 1, 2, 3)

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
        self.err_output.map_read()
        self.err_input.map_invalidate()
        self.err_input.mem[...] = self.err_output.mem.reshape(self.output.shape).transpose(self.PERM)

    def b200_run(self):
        ei = self.err_input.devmem.reshape(self.input.shape)
        ei.copy_(self.err_output.devmem.reshape(self.output.shape).permute(*self.PERM))
        self.err_input.mark_device_written()


class GDChannelSplitter(_GDPermute):
    """Backward pass of ChannelSplitter."""
    hide_from_registry = False
    MAPPING = {"channel_splitter"}
    PERM = (0, 2, 3, 1)


class GDChannelMerger(_GDPermute):
    """Backward pass of ChannelMerger."""
    hide_from_registry = False
    MAPPING = {"channel_merger"}
    PERM = (0, 3, 1, 2)
