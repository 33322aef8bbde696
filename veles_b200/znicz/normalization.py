"""Local response normalisation across channels (``norm`` layer type; params ``alpha, beta, n, k``
— ``docs/source/manualrst_veles_workflow_parameters.rst:63-64, 497, 538-541``).
``y_c = x_c * (k + alpha * Σ_{c' in window(c, n)} x_{c'}²)^(-beta)`` (AlexNet LRN)."""
from __future__ import annotations

import numpy

from .nn_units import ForwardBase, GDBase, act_deriv_numpy


def _window_sums(sq, n):
    """Σ over a centred window of n channels (last axis), zero padded."""
    half = n // 2
    C = sq.shape[-1]
    pad = numpy.zeros(sq.shape[:-1] + (C + 2 * half,), dtype=sq.dtype)
    pad[..., half:half + C] = sq
    cs = numpy.concatenate([numpy.zeros(sq.shape[:-1] + (1,), dtype=sq.dtype), numpy.cumsum(pad, axis=-1)], axis=-1)
    return cs[..., 2 * half + 1:2 * half + 1 + C] - cs[..., 0:C]


class LRNormalizerForward(ForwardBase):
    """Local response normalisation across channels: y = x (k + alpha sum_{window} x^2)^-beta."""
    hide_from_registry = False
    MAPPING = {"norm"}
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0021"

    def __init__(self, workflow, **kwargs):
        super(LRNormalizerForward, self).__init__(workflow, **kwargs)
        self.alpha = float(kwargs.get("alpha", 0.0001))
        self.beta = float(kwargs.get("beta", 0.75))
        self.k = float(kwargs.get("k", 2))
        self.n = int(kwargs.get("n", 5))
        self.exports = ["alpha", "beta", "k", "n"]

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        self.alloc_output(self.input.shape)
        if self.on_device:
            self.output.dev_dtype = self.input.devmem.dtype
            self.output.initialize(self.device_)
        self._backend_init_()
        return None

    def package_export(self):
        return {"alpha": self.alpha, "beta": self.beta, "k": self.k, "n": self.n}

    def numpy_run(self):
        self.input.map_read()
        x = self.input.mem.astype(numpy.float64)
        s = self.k + self.alpha * _window_sums(x * x, self.n)
        self.output.map_invalidate()
        self.output.mem[...] = x * s ** (-self.beta)

    def b200_run(self):
        from ..ops import kernels
        x = self.input.devmem
        kernels.lrn_forward(x if x.is_contiguous() else x.contiguous(), self.output.devmem, self.n, self.alpha,
                            self.beta, self.k)
        self.output.mark_device_written()


class LRNormalizerBackward(GDBase):
    """Backward pass of local response normalisation (can also apply the previous layer's activation derivative
    and leave that layer's bias-gradient sums)."""
    hide_from_registry = False
    MAPPING = {"norm"}
    CAN_FUSE_PREV_DERIVATIVE = True

    def __init__(self, workflow, **kwargs):
        super(LRNormalizerBackward, self).__init__(workflow, **kwargs)
        self.demand("input")
        self.pool_gd = None      # backward unit of the max pooling layer that follows (StandardWorkflow fusion plan)
        # GD unit of the convolution in front of this LRN whose bias gradient (= column sums of this unit's err_input)
        # is accumulated by the LRN backward kernel itself (StandardWorkflow._plan_bias_sum_fusion)
        self.bias_sum_for = None

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
        self.input.map_read()
        self.err_output.map_read()
        x = self.input.mem.astype(numpy.float64)
        dy = self.err_output.mem.reshape(x.shape).astype(numpy.float64)
        s = f.k + f.alpha * _window_sums(x * x, f.n)
        t = dy * x * s ** (-f.beta - 1.0)
        dx = dy * s ** (-f.beta) - 2.0 * f.alpha * f.beta * x * _window_sums(t, f.n)
        if self.prev_activation:
            dx = dx * act_deriv_numpy(self.prev_activation, x)
        self.err_input.map_invalidate()
        self.err_input.mem[...] = dx

    def b200_run(self):
        from ..ops import kernels
        f = self.forward
        x = self.input.devmem
        ei = self.err_input.devmem
        pg = self.pool_gd
        if pg is not None and getattr(pg, "handed_over_", False):
            # the pooling backward unit skipped its pass: scatter its err_output by the argmax bytes and apply the
            # LRN derivative in one kernel
            pf = pg.forward
            x4 = x.reshape(pf.in_shape4)
            if not kernels.lrn_pool_backward(x4, pg.err_output.devmem.reshape(pf.out_shape4), pf.offs8_,
                                             ei.reshape(pf.in_shape4), pf.ky, pf.kx, pf.sliding[1], pf.sliding[0],
                                             f.n, f.alpha, f.beta, f.k, deriv=self.prev_activation):
                raise RuntimeError("%s: fused pooling + LRN backward refused a geometry its planner accepted" % self.name)
            self.err_input.mark_device_written()
            if self.bias_sum_for is not None:
                self.bias_sum_for.bias_partials_ = None
            return
        eo = self.err_output.devmem
        if eo.dtype != x.dtype or not eo.is_contiguous():
            eo = eo.contiguous().to(x.dtype)
        tgt = self.bias_sum_for
        done = None
        if tgt is not None:
            import torch
            C = x.shape[-1]
            if getattr(self, "bias_partials_", None) is None:
                self.bias_partials_ = torch.zeros(max(1, self.device_.device_info.sm_count) * 4, C, dtype=torch.float32, device=x.device)
            done = kernels.lrn_backward(x, eo.view(x.shape), ei, f.n, f.alpha, f.beta, f.k,
                                        deriv=self.prev_activation, partials=self.bias_partials_)
            # the convolution's GD unit folds the rows (and applies its bias step) instead of summing err_output
            tgt.bias_partials_ = (self.bias_partials_, done) if done is not None else None
        if done is None:
            kernels.lrn_backward(x, eo.view(x.shape), ei, f.n, f.alpha, f.beta, f.k, deriv=self.prev_activation)
        self.err_input.mark_device_written()
