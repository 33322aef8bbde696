"""Backward units of the pooling layers: ``GDMaxPooling``, ``GDMaxAbsPooling``, ``GDAvgPooling``
and the stochastic variants (``docs/source/manualrst_veles_workflow_parameters.rst:470-504``).
Max-like: ``err_input[input_offset] += err_output``; avg: the error is spread uniformly over
the (clipped) window.  Device kernel is in gather form (deterministic, no atomics) and can
fold in the derivative of the activation that produced the pooling input.
"""
from __future__ import annotations

import numpy

from .nn_units import GDBase, act_deriv_numpy


class GDPooling(GDBase):
    hide_from_registry = True
    KIND = "max"
    FORWARD_ATTRS = ()
    CAN_FUSE_PREV_DERIVATIVE = True

    def __init__(self, workflow, **kwargs):
        super(GDPooling, self).__init__(workflow, **kwargs)
        self.demand("input")
        self.lrn_consumer = None     # LRN backward unit that consumes err_input (StandardWorkflow fusion plan)

    def init_unpickled(self):
        super(GDPooling, self).init_unpickled()
        self.handed_over_ = False

    def hands_over_to_lrn(self):
        """True when this run's work is done by the LRN backward unit's fused kernel: 3x3 / stride 2 max pooling with
        byte argmax on a bf16 device tensor of at most 256 channels."""
        from ..config import root, get
        f = self.forward
        c = self.lrn_consumer
        # opt-in: measured on AlexNet the per-pixel gather makes the (instruction-bound) LRN backward kernel slower
        # than the two separate kernels (step 2.46 vs 2.33 ms) - see DESIGN.md section 8
        if c is None or not self.on_device or not get(root.common.engine.fuse_lrn_pool, False):
            return False
        if getattr(f, "offs8_", None) is None or self.prev_activation:
            return False
        import torch
        eo = self.err_output.devmem
        C = f.in_shape4[3]
        return (f.ky == 3 and f.kx == 3 and tuple(f.sliding) == (2, 2) and C % 8 == 0 and C <= 256
                and 3 <= getattr(c.forward, "n", 0) <= 9 and eo is not None and eo.dtype == torch.bfloat16
                and eo.is_contiguous() and c.input.devmem is not None and c.input.devmem.dtype == torch.bfloat16
                and c.input.devmem.is_contiguous())

    def initialize(self, device=None, **kwargs):
        if self.err_output is None or not self.err_output or self.input is None or not self.input:
            return True
        if self.forward is None:
            raise ValueError("%s needs forward=<pooling unit>" % type(self).__name__)
        self.device = device
        self.alloc_err_input()
        self._backend_init_()
        return None

    def numpy_run(self):
        f = self.forward
        self.err_output.map_read()
        eo = self.err_output.mem.reshape(f.out_shape4).astype(numpy.float32)
        B, H, W, C = f.in_shape4
        ei = numpy.zeros((B, H, W, C), dtype=numpy.float32)
        if f.KIND == "avg":
            for oh, ow, h0, h1, w0, w1 in f._windows():
                ei[:, h0:h1, w0:w1, :] += (eo[:, oh, ow, :] / ((h1 - h0) * (w1 - w0)))[:, None, None, :]
        else:
            f.input_offset.map_read()
            numpy.add.at(ei.reshape(-1), f.input_offset.mem.reshape(-1), eo.reshape(-1))
        if self.prev_activation:
            self.input.map_read()
            ei *= act_deriv_numpy(self.prev_activation, self.input.mem.reshape(ei.shape).astype(numpy.float32))
        self.err_input.map_invalidate()
        self.err_input.mem[...] = ei.reshape(self.err_input.shape)

    def b200_init(self):
        dev = self.device_
        self.err_input.dev_dtype = dev.compute_dtype
        self.err_input.initialize(dev)

    def b200_run(self):
        from ..ops import kernels
        f = self.forward
        self.err_output.unmap()
        self.handed_over_ = self.hands_over_to_lrn()
        if self.handed_over_:
            return       # the LRN backward unit gathers straight from err_output + argmax bytes
        eo = self.err_output.devmem
        if eo.dtype != self.err_input.devmem.dtype or not eo.is_contiguous():
            if getattr(self, "_eo_stage_", None) is None:
                import torch
                self._eo_stage_ = torch.empty(f.out_shape4, dtype=self.err_input.devmem.dtype, device=eo.device)
            B = f.out_shape4[0]
            kernels.copy_2d(eo.reshape(B, -1) if eo.dim() != 2 else eo, self._eo_stage_.view(B, -1))
            eo = self._eo_stage_
        dy = eo.reshape(f.out_shape4)
        dx = self.err_input.devmem.reshape(f.in_shape4)
        z = self.input.devmem.reshape(f.in_shape4) if self.prev_activation else None
        if getattr(f, "offs8_", None) is not None and kernels.pool_backward_local(
                kernels.POOL_KINDS[f.KIND], dy, f.offs8_, dx, f.ky, f.kx, f.sliding[1], f.sliding[0], z=z,
                deriv=self.prev_activation):
            self.err_input.mark_device_written()
            return
        if f.KIND != "avg" and hasattr(f, "ensure_input_offset"):
            f.ensure_input_offset()
        kernels.pool_backward(kernels.POOL_KINDS[f.KIND], dy, f.input_offset.devmem if f.KIND != "avg" else None,
                              dx, f.ky, f.kx, f.sliding[1], f.sliding[0], z=z, deriv=self.prev_activation)
        self.err_input.mark_device_written()


class GDMaxPooling(GDPooling):
    """Backward pass of max (and stochastic) pooling: routes each gradient to its window's recorded position."""
    hide_from_registry = False
    MAPPING = {"max_pooling", "stochastic_pooling", "stochastic_abs_pooling"}


class GDMaxAbsPooling(GDPooling):
    """Backward pass of max-abs pooling."""
    hide_from_registry = False
    MAPPING = {"maxabs_pooling"}


class GDAvgPooling(GDPooling):
    """Backward pass of average pooling."""
    hide_from_registry = False
    MAPPING = {"avg_pooling"}


class GDPoolDepool(GDBase):
    """Backward of the pool-depool pair: pass the error only through the chosen elements."""
    hide_from_registry = False
    MAPPING = {"stochastic_pool_depool", "stochastic_abs_pool_depool"}

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
        f.input_offset.map_read()
        eo = self.err_output.mem.reshape(-1)
        ei = numpy.zeros(eo.shape, dtype=numpy.float32)
        idx = f.input_offset.mem.reshape(-1)
        ei[idx] = eo[idx]
        self.err_input.map_invalidate()
        self.err_input.mem[...] = ei.reshape(self.err_input.shape)

    def b200_run(self):
        f = self.forward
        eo = self.err_output.devmem.reshape(-1)
        ei = self.err_input.devmem.view(-1)
        ei.zero_()
        idx = f.input_offset.devmem.view(-1).long()
        ei[idx] = eo[idx].to(ei.dtype)
        self.err_input.mark_device_written()
