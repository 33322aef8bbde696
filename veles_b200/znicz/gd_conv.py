"""Gradient descent for convolutional layers: ``GradientDescentConv``, ``GDTanhConv``,
``GDRELUConv``, ``GDStrictRELUConv``, ``GDSigmoidConv``
(``docs/source/manualrst_veles_workflow_parameters.rst:470-504``).

Device path per layer: using the column matrix kept by the forward unit,
  dW_g  += err_output[:, g]ᵀ · cols[:, g]      tcgen05 GEMM, MN-major x MN-major, split-K, fp32 atomics
  db     = Σ_pixels err_output                 deterministic two-phase column sum
  dcols[:, g] = err_output[:, g] · W_g         tcgen05 GEMM (overwrites cols: it is dead after dW)
  err_input = col2im(dcols) ∘ f'_prev(input)   gather-form col2im with the fused derivative
"""
from __future__ import annotations

import numpy

from .conv import numpy_im2col, numpy_col2im
from .nn_units import GradientDescentBase, act_deriv_numpy


class GradientDescentConv(GradientDescentBase):
    """Backward pass and update of a convolutional layer: err_input and weight-gradient as implicit GEMMs, bias
    sums, Znicz update."""
    hide_from_registry = False
    MAPPING = {"conv"}
    ACTIVATION = 0
    CAN_FUSE_PREV_DERIVATIVE = True

    def __init__(self, workflow, **kwargs):
        super(GradientDescentConv, self).__init__(workflow, **kwargs)
        self.demand("input", "output", "weights")

    def initialize(self, device=None, **kwargs):
        if self.err_output is None or not self.err_output or self.input is None or not self.input:
            return True
        if self.forward is None:
            raise ValueError("GradientDescentConv needs forward=<Conv unit> (geometry and column matrix)")
        self.device = device
        self.alloc_err_input()
        self._ensure_state(self.weights, self.gradient_weights_with_moment, self.accumulated_gradient_weights)
        if self.bias is not None and self.bias:
            self._ensure_state(self.bias, self.gradient_bias_with_moment, self.accumulated_gradient_bias)
        self._backend_init_()
        return None

    def numpy_run(self):
        f = self.forward
        self.err_output.map_read()
        self.input.map_read()
        self.weights.map_write()
        nk = f.n_kernels
        eo = self.err_output.mem.reshape(-1, nk).astype(numpy.float32)
        if self.ACTIVATION and not self.derivative_applied:
            self.output.map_read()
            eo = eo * act_deriv_numpy(self.ACTIVATION, self.output.mem.reshape(eo.shape))
            self.err_output.map_write()
            self.err_output.mem[...] = eo.reshape(self.err_output.shape)
        x = self.input.mem.reshape(f.in_shape4).astype(numpy.float32)
        pl, pt, pr, pb = f.padding
        sx, sy = f.sliding
        G = f.grouping
        cols, OH, OW = numpy_im2col(x, f.ky, f.kx, sy, sx, pl, pt, pr, pb, G)
        w = self.weights.mem
        nkg = nk // G
        kg = cols.shape[1] // G
        gw = numpy.empty_like(w)
        dcols = numpy.empty_like(cols) if self.need_err_input else None
        for gi in range(G):
            e = eo[:, gi * nkg:(gi + 1) * nkg]
            gw[gi * nkg:(gi + 1) * nkg] = e.T.dot(cols[:, gi * kg:(gi + 1) * kg])
            if self.need_err_input:
                dcols[:, gi * kg:(gi + 1) * kg] = e.dot(w[gi * nkg:(gi + 1) * nkg])
        if self.need_err_input:
            ei = numpy_col2im(dcols, x.shape, f.ky, f.kx, sy, sx, pl, pt, pr, pb, G)
            if self.prev_activation:
                ei = ei * act_deriv_numpy(self.prev_activation, x)
            self.err_input.map_invalidate()
            self.err_input.mem[...] = ei.reshape(self.err_input.shape)
        gw, = self.dp_reduce_host(gw)
        self.gradient_weights.reset(gw.astype(numpy.float32))
        if self.apply_gradient:
            self.numpy_update(w, gw, self.gradient_weights_with_moment.mem,
                              self.accumulated_gradient_weights.mem if self.accumulated_gradient_weights else None,
                              self.learning_rate, self.weights_decay, self.l1_vs_l2, self.gradient_moment)
        if self.bias is not None and self.bias:
            self.bias.map_write()
            gb = eo.sum(axis=0)
            gb, = self.dp_reduce_host(gb)
            self.gradient_bias.reset(gb.astype(numpy.float32))
            if self.apply_gradient:
                self.numpy_update(self.bias.mem, gb, self.gradient_bias_with_moment.mem,
                                  self.accumulated_gradient_bias.mem if self.accumulated_gradient_bias else None,
                                  self.learning_rate_bias, self.weights_decay_bias, self.l1_vs_l2_bias,
                                  self.gradient_moment_bias)

    def b200_init(self):
        import torch
        from ..ops import kernels
        dev = self.device_
        self.place_state()
        if self.need_err_input:
            self.err_input.dev_dtype = dev.compute_dtype
            self.err_input.initialize(dev)
        nk = self.forward.n_kernels
        rows = self.err_output.size // nk
        self.col_scratch_ = torch.empty(kernels.col_sum_scratch_elems(rows, nk), dtype=torch.float32,
                                        device=dev.torch_device)

    def b200_run(self):
        import torch
        from ..ops import kernels
        f = self.forward
        nk = f.n_kernels
        self.err_output.unmap()
        eo = self.err_output.devmem
        eo2 = eo.reshape(-1, nk) if eo.dim() != 2 or eo.shape[1] != nk else eo
        if self.ACTIVATION and not self.derivative_applied:
            y = self.output.devmem
            kernels.activation_backward(self.ACTIVATION, eo2, None, y.reshape(-1, nk), eo2)
        if eo2.dtype != torch.bfloat16 or eo2.stride(0) % 8 != 0:
            # TMA operand rules (bf16, row pitch a multiple of 16 bytes): stage odd kernel counts into a pitched copy
            if getattr(self, "_eo_stage_", None) is None:
                self._eo_stage_ = kernels.pitched_empty(eo2.shape[0], nk, torch.bfloat16, eo2.device)
            kernels.copy_2d(eo2, self._eo_stage_)
            eo2 = self._eo_stage_
        G = f.grouping
        nkg = nk // G
        w_bf16 = f.weights_bf16_
        kg = w_bf16.shape[1]
        gw_t, gb_t = self.grad_tensors()
        gw = gw_t.view(nk, kg)
        g = f.geometry
        x4 = f.device_input4() if f._x_stage4_ is None else f._x_stage4_
        Cg = x4.shape[3] // G
        if f.s2d_ is not None:
            # the forward pass left the space-to-depth input in place; gradient in s2d column order, then folded
            # into the Znicz-ordered gradient (and the scratch cleared for the next step)
            sd = f.s2d_
            kernels.conv_weight_grad(sd["x"], eo2, sd["gw"], sd["geom"], accumulate=True, split_k=0)
            kernels.gather_cols(sd["gw"], gw, sd["from_s2d"], accumulate=True)
            sd["gw"].zero_()
        elif f.implicit_:
            for gi in range(G):
                kernels.conv_weight_grad(x4[..., gi * Cg:(gi + 1) * Cg] if G > 1 else x4,
                                         eo2[:, gi * nkg:(gi + 1) * nkg] if G > 1 else eo2,
                                         gw[gi * nkg:(gi + 1) * nkg], g, accumulate=True, split_k=0)
        else:
            cols = f.cols_
            for gi in range(G):
                e = eo2[:, gi * nkg:(gi + 1) * nkg] if G > 1 else eo2
                kernels.gemm(e, cols[:, gi * kg:(gi + 1) * kg], gw[gi * nkg:(gi + 1) * nkg], a_mn=True, b_mn=True,
                             accumulate=True, split_k=0)
        self.gradient_weights.mark_device_written()
        if self.bias is not None and self.bias:
            if self.__dict__.get("bias_partials_") is not None:
                # the unit that produced err_output left per-block column sums of it
                self.device_bias_from_partials(gb_t, *self.bias_partials_)
                self.bias_partials_ = None
            else:
                self.device_bias_sum(eo2, gb_t)
        if self.need_err_input:
            ei = self.err_input.devmem
            ei4 = ei.reshape(f.in_shape4) if tuple(ei.shape) != tuple(f.in_shape4) else ei
            z = x4 if self.prev_activation else None
            if f.implicit_ and g["sx"] == 1 and g["sy"] == 1:
                # conv-transpose as one more implicit GEMM over dY; writes NHWC err_input directly
                eo4 = eo2.view(x4.shape[0], g["OH"], g["OW"], nk)
                ei2 = ei4.view(-1, x4.shape[3])
                z2 = z.view(-1, x4.shape[3]) if z is not None else None
                if kernels.grouped_conv_ok(x4.shape[3], nk, G):
                    kernels.conv_err_input_grouped(eo4, w_bf16, ei2, g, G, z=z2, deriv=self.prev_activation)
                    self.err_input.mark_device_written()
                    self.device_update()
                    return
                halo = kernels.conv_halo_preferred(Cg, g)
                for gi in range(G):
                    args = (eo4[..., gi * nkg:(gi + 1) * nkg] if G > 1 else eo4, w_bf16[gi * nkg:(gi + 1) * nkg],
                            ei2[:, gi * Cg:(gi + 1) * Cg] if G > 1 else ei2, g)
                    kw = dict(z=(z2[:, gi * Cg:(gi + 1) * Cg] if G > 1 else z2) if z2 is not None else None,
                              deriv=self.prev_activation)
                    if not (halo and kernels.conv_err_input_halo(*args, **kw)):
                        kernels.conv_err_input(*args, **kw)
            else:
                cols = f.ensure_cols()   # (also the err_input scratch of strided / space-to-depth layers)
                for gi in range(G):
                    e = eo2[:, gi * nkg:(gi + 1) * nkg] if G > 1 else eo2
                    kernels.gemm(e, w_bf16[gi * nkg:(gi + 1) * nkg], cols[:, gi * kg:(gi + 1) * kg], b_mn=True)
                kernels.col2im(cols, ei4, g, z=z, deriv=self.prev_activation)
            self.err_input.mark_device_written()
        self.device_update()


class GDTanhConv(GradientDescentConv):
    """GradientDescentConv for ConvTanh."""
    MAPPING = {"conv_tanh"}
    ACTIVATION = 1


class GDRELUConv(GradientDescentConv):
    """GradientDescentConv for ConvRELU."""
    MAPPING = {"conv_relu"}
    ACTIVATION = 3


class GDStrictRELUConv(GradientDescentConv):
    """GradientDescentConv for ConvStrictRELU."""
    MAPPING = {"conv_str"}
    ACTIVATION = 4


class GDSigmoidConv(GradientDescentConv):
    """GradientDescentConv for ConvSigmoid."""
    MAPPING = {"conv_sigmoid"}
    ACTIVATION = 2
