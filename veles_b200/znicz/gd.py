"""Gradient descent units for fully connected layers.

Types (``docs/source/manualrst_veles_workflow_parameters.rst:470-504, 556-578``):
``GradientDescent`` (linear), ``GDTanh``, ``GDRELU``, ``GDStrictRELU``, ``GDSigmoid``,
``GDSoftmax`` (its err_output already is dL/d(pre-softmax)).

Device path per layer:
  dX  = (err_output · W) ∘ f'_prev(input)      tcgen05 GEMM, A K-major, B MN-major, derivative epilogue
  dW  = err_outputᵀ · input                     tcgen05 GEMM, both operands MN-major, fp32 out
  db  = Σ_batch err_output                      two-phase column sum (deterministic)
  W,b ← fused GD step (+ bf16 shadow)           one kernel, or the fused all-reduce+GD kernel in DP mode
"""
from __future__ import annotations

import numpy

from ..memory import Array
from .nn_units import GradientDescentBase, act_deriv_numpy


class GradientDescent(GradientDescentBase):
    """Backward pass and update of a fully connected linear layer: err_input GEMM, weight-gradient GEMM, bias
    sums, Znicz momentum / decay update (fused on the device)."""
    hide_from_registry = False
    MAPPING = {"all2all"}
    ACTIVATION = 0
    CAN_FUSE_PREV_DERIVATIVE = True

    def __init__(self, workflow, **kwargs):
        super(GradientDescent, self).__init__(workflow, **kwargs)
        self.demand("input", "output", "weights")

    def initialize(self, device=None, **kwargs):
        if self.err_output is None or not self.err_output or self.input is None or not self.input:
            return True
        self.device = device
        if self.need_err_input and (not self.err_input or self.err_input.shape != self.input.shape):
            self.err_input.reset(numpy.zeros(self.input.shape, dtype=numpy.float32))
        if self.weights:
            self._ensure_state(self.weights, self.gradient_weights_with_moment, self.accumulated_gradient_weights)
        if self.bias is not None and self.bias:
            self._ensure_state(self.bias, self.gradient_bias_with_moment, self.accumulated_gradient_bias)
        self._backend_init_()
        return None

    # -- numpy oracle ---------------------------------------------------------------------------------------
    def numpy_err_output(self):
        self.err_output.map_read()
        eo = self.err_output.mem.reshape(self.err_output.shape[0], -1).astype(numpy.float32)
        if self.ACTIVATION and not self.derivative_applied:
            self.output.map_read()
            eo = eo * act_deriv_numpy(self.ACTIVATION, self.output.mem.reshape(eo.shape))
            # keep Znicz behaviour: err_output is updated in place
            self.err_output.map_write()
            self.err_output.mem[...] = eo.reshape(self.err_output.shape)
        return eo

    def numpy_run(self):
        eo = self.numpy_err_output()
        self.input.map_read()
        self.weights.map_write()
        x = self.input.mem.reshape(self.input.shape[0], -1).astype(numpy.float32)
        w = self.weights.mem
        transposed = bool(getattr(self.forward, "weights_transposed", False)) if self.forward is not None \
            else (w.shape[0] != eo.shape[1])
        if self.need_err_input:
            self.err_input.map_invalidate()
            ei = eo.dot(w.T if transposed else w)
            if self.prev_activation:
                ei = ei * act_deriv_numpy(self.prev_activation, x)
            self.err_input.mem[...] = ei.reshape(self.err_input.shape)
        gw = eo.T.dot(x)
        if transposed:
            gw = gw.T
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

    # -- device ------------------------------------------------------------------------------------------------
    def b200_init(self):
        import torch
        from ..ops import kernels
        dev = self.device_
        self.place_state()
        if self.need_err_input:
            self.err_input.dev_dtype = dev.compute_dtype
            self.err_input._pitched = True
            self.err_input.initialize(dev)
        n_out = self.err_output.size // self.err_output.shape[0]
        self.col_scratch_ = torch.empty(kernels.col_sum_scratch_elems(self.err_output.shape[0], n_out),
                                        dtype=torch.float32, device=dev.torch_device)
        self._eo_stage_ = None
        # dW = dY^T X either accumulates into a zeroed buffer (split-K over the batch) or, when the scheduler would
        # not split, is plainly stored: the latter skips the read-modify-write at L2 and the re-zeroing pass.
        n_in = self.input.size // self.input.shape[0]
        # err_input = dY W with few M tiles: same two-phase split-K as the forward pass
        self.ei_acc_ = None
        if (getattr(self.forward, "TWO_PHASE_SPLIT_K", False) and self.need_err_input and not self.tf32_mode()
                and n_in % 4 == 0
                and kernels.gemm_plan(self.err_output.shape[0], n_in, n_out, b_mn=True)[1] >= 2):
            self.ei_acc_ = torch.zeros(self.err_output.shape[0], n_in, dtype=torch.float32, device=dev.torch_device)
        self.transposed_ = bool(getattr(self.forward, "weights_transposed", False))
        dw_m, dw_n = (n_in, n_out) if self.transposed_ else (n_out, n_in)
        self.dw_store_ = (self.dp_group_ is None and self.apply_gradient and not self.tf32_mode() and
                          kernels.gemm_plan(dw_m, dw_n, self.err_output.shape[0], b_mn=True)[1] == 1)
        # Optionally (root.common.engine.fuse_gd_update) the GD step runs in that GEMM's epilogue: the update reads w
        # and momentum at the tile's position and the gradient never reaches HBM (single GPU, momentum solver).
        # Measured on AlexNet's fc6/fc7 it only breaks even (+236 us in the GEMMs, -193 us of update kernels): the
        # separate update already finds the freshly written gradient in the 126 MB L2 -> off by default.
        from ..config import root, get
        self.fused_gd_ = bool(self.dw_store_ and self.solver_code == 0 and not self.accumulate_gradient and
                              not self.factor_ortho and
                              not self.transposed_ and
                              self.forward is not None and getattr(self.forward, "weights_bf16_", None) is not None and
                              get(root.common.engine.fuse_gd_update, False))

    def device_err_output(self):
        """bf16 2-D err_output with own activation derivative applied (unless fused upstream)."""
        import torch
        from ..ops import kernels
        eo = self.err_output.devmem
        eo2 = eo if eo.dim() == 2 else eo.reshape(eo.shape[0], -1)
        if self.tf32_mode():     # fp32 err_output straight into the tf32 GEMMs
            if eo2.dtype != torch.float32 or eo2.stride(0) % 4 != 0:
                eo2 = eo2.float().contiguous()
            if self.ACTIVATION and not self.derivative_applied:
                y = self.output.devmem
                y2 = (y if y.dim() == 2 else y.reshape(y.shape[0], -1))
                tmp = eo2.contiguous()
                kernels.activation_backward(self.ACTIVATION, tmp, None, y2.contiguous(), tmp)
                eo2 = tmp
            return eo2
        bad_layout = eo2.dtype != torch.bfloat16 or eo2.stride(0) % 8 != 0
        if bad_layout:
            if self._eo_stage_ is None:
                self._eo_stage_ = kernels.pitched_empty(eo2.shape[0], eo2.shape[1], torch.bfloat16, eo2.device)
            kernels.copy_2d(eo2, self._eo_stage_)
            eo2 = self._eo_stage_
        if self.ACTIVATION and not self.derivative_applied:
            y = self.output.devmem
            y2 = y if y.dim() == 2 else y.reshape(y.shape[0], -1)
            if eo2.is_contiguous() and y2.is_contiguous():
                kernels.activation_backward(self.ACTIVATION, eo2, None, y2, eo2)
            else:  # pitched rows: go row-block wise through a contiguous temp
                tmp = eo2.contiguous()
                kernels.activation_backward(self.ACTIVATION, tmp, None, y2.contiguous(), tmp)
                kernels.copy_2d(tmp, eo2)
        return eo2

    def tf32_mode(self):
        return bool(self.forward is not None and getattr(self.forward, "tf32_", False))

    def b200_run_tf32(self):
        """The three GEMMs of the layer on fp32 tensors through tcgen05 kind::tf32."""
        from ..ops import kernels
        fwd = self.forward
        eo = self.device_err_output()
        x = fwd.device_input_2d()
        w = fwd.weights.devmem.view(fwd.weights_bf16_.shape)
        if self.need_err_input:
            ei = self.err_input.devmem
            ei2 = ei if ei.dim() == 2 else ei.reshape(ei.shape[0], -1)
            kernels.gemm_tf32(eo, w, ei2, b_mn=True, z=x if self.prev_activation else None, deriv=self.prev_activation)
            self.err_input.mark_device_written()
        gw_t, gb_t = self.grad_tensors()
        gw = gw_t.view(w.shape[0], -1)
        kernels.gemm_tf32(eo, x, gw, a_mn=True, b_mn=True, accumulate=True, split_k=0)
        if self.bias is not None and self.bias:
            self.device_bias_sum(eo, gb_t)
        self.gradient_weights.mark_device_written()
        self.device_update()

    def b200_run(self):
        from ..ops import kernels
        self.err_output.unmap()
        if self.tf32_mode():
            return self.b200_run_tf32()
        eo = self.device_err_output()
        fwd = self.forward
        # the forward unit staged (cast/pitched) its input this iteration: reuse it
        x = fwd._x_stage_ if (fwd is not None and getattr(fwd, "_x_stage_", None) is not None) \
            else self._input_2d()
        w_bf16 = fwd.weights_bf16_ if fwd is not None else self._own_shadow()
        if self.need_err_input:
            ei = self.err_input.devmem
            ei2 = ei if ei.dim() == 2 else ei.reshape(ei.shape[0], -1)
            peers = getattr(self.forward, "peer_weights_", None)
            if self.transposed_:
                # weights are (n_in, n_out): dY . W^T reads them K-major, like a forward GEMM
                kernels.gemm(eo, w_bf16, ei2, z=x if self.prev_activation else None, deriv=self.prev_activation)
            elif self.ei_acc_ is not None and peers is None:
                kernels.gemm(eo, w_bf16, self.ei_acc_, b_mn=True, accumulate=True, split_k=0)
                kernels.bias_act(self.ei_acc_, ei2, z=x if self.prev_activation else None, deriv=self.prev_activation)
            else:
                kernels.gemm(eo, w_bf16, ei2, b_mn=True, z=x if self.prev_activation else None,
                             deriv=self.prev_activation, b_peers=peers)
            self.err_input.mark_device_written()
        gw_t, gb_t = self.grad_tensors()
        gw = gw_t.view(w_bf16.shape[0], -1)
        # (a data-parallel group attaches after initialize: its gradients must be exchanged, never fused away)
        fused = getattr(self, "fused_gd_", False) and self.apply_gradient and self.dp_group_ is None
        if fused:
            w2d = self.weights.devmem.view(w_bf16.shape[0], -1)
            kernels.gemm_fused_gd(eo, x, w2d, self.gradient_weights_with_moment.devmem.view(w2d.shape), w_bf16,
                                  lr=self.eff(self.learning_rate, 0)[0], wd=self.weights_decay, l1_vs_l2=self.l1_vs_l2,
                                  moment=self.eff(0, self.gradient_moment)[1])
            self.weights.mark_device_written()
            self.gradient_weights_with_moment.mark_device_written()
        elif self.dw_store_:
            ga, gb = (x, eo) if self.transposed_ else (eo, x)        # transposed weights: dW = X^T . dY
            kernels.gemm(ga, gb, gw, a_mn=True, b_mn=True, accumulate=False, split_k=1)
        else:
            ga, gb = (x, eo) if self.transposed_ else (eo, x)
            kernels.gemm(ga, gb, gw, a_mn=True, b_mn=True, accumulate=True, split_k=0)
        if self.bias is not None and self.bias:
            self.device_bias_sum(eo, gb_t)
        if not fused:
            self.gradient_weights.mark_device_written()
        self.device_update(skip_weights=fused)

    def _input_2d(self):
        import torch
        from ..ops import kernels
        x = self.input.devmem
        x2 = x if x.dim() == 2 else x.reshape(x.shape[0], -1)
        if x2.dtype != torch.bfloat16 or x2.stride(0) % 8 != 0:
            if getattr(self, "_x_stage_", None) is None:
                self._x_stage_ = kernels.pitched_empty(x2.shape[0], x2.shape[1], torch.bfloat16, x2.device)
            kernels.copy_2d(x2, self._x_stage_)
            return self._x_stage_
        return x2

    def _own_shadow(self):
        import torch
        from ..ops import kernels
        if getattr(self, "_shadow_", None) is None:
            n_out = self.weights.shape[0]
            n_in = self.weights.size // n_out
            self._shadow_ = kernels.pitched_empty(n_out, n_in, torch.bfloat16, self.device_.torch_device)
        kernels.copy_2d(self.weights.devmem.view(self._shadow_.shape[0], -1), self._shadow_)
        return self._shadow_


class GDTanh(GradientDescent):
    """GradientDescent for All2AllTanh."""
    MAPPING = {"all2all_tanh"}
    ACTIVATION = 1


class GDRELU(GradientDescent):
    """GradientDescent for All2AllRELU."""
    MAPPING = {"all2all_relu"}
    ACTIVATION = 3


class GDStrictRELU(GradientDescent):
    """GradientDescent for All2AllStrictRELU."""
    MAPPING = {"all2all_str"}
    ACTIVATION = 4


class GDSigmoid(GradientDescent):
    """GradientDescent for All2AllSigmoid."""
    MAPPING = {"all2all_sigmoid"}
    ACTIVATION = 2


class GDSoftmax(GradientDescent):
    """err_output from EvaluatorSoftmax is already the gradient w.r.t. the pre-softmax sums."""
    MAPPING = {"softmax"}
    ACTIVATION = 0
