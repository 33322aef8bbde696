"""Recurrent layers: Elman RNN and LSTM, with back-propagation through time.

The reference lists both among Znicz's algorithms ("all necessary units developed for CUDA, OpenCL and numpy",
``docs/source/manualrst_veles_algorithms.rst:108-135``) but ships no source for them, so the semantics are pinned by
the numpy oracle in this file:

    RNN :  h_t = tanh([x_t | h_{t-1}] . W^T + b)                                  W: (H, D + H)
    LSTM:  (i, f, g, o) = [x_t | h_{t-1}] . W^T + b     split in four blocks of H  W: (4H, D + H), gate order i f g o
           c_t = sigm(f) * c_{t-1} + sigm(i) * tanh(g);   h_t = sigm(o) * tanh(c_t)

``input`` is ``[batch, T, D]`` (or ``[batch, T * D]`` with ``sequence_length=T``); ``output`` is ``[batch, T, H]``, or
the last state ``[batch, H]`` with ``return_sequences=False``.  h_0 = c_0 = 0 for every minibatch.

Device path (B200): the concatenated operand ``[x_t | h_{t-1}]`` lives time-major in ONE bf16 buffer ``xh``
(``[T * batch, D + H]``), so that
  * a forward step is one tcgen05 GEMM (bias in the epilogue, fp32 pre-activations) + one cell kernel
    (``csrc/recurrent.cu``) that writes h_t directly into step t+1's slot of ``xh`` and into ``output``;
  * a backward step is one cell kernel (dZ_t) + one GEMM ``dZ_t . W`` that yields dx_t and dh_{t-1} together;
  * the weight gradient is a single GEMM ``dZ^T . xh`` over all T * batch rows and the bias gradient a single column
    sum, after the loop - the same two launches a fully connected layer needs, whatever T is.
The weights are one ``(G * H, D + H)`` matrix, so the fused GD update, the bf16 shadow, snapshots and the data-parallel
all-reduce+GD kernel treat a recurrent layer exactly like a fully connected one.
"""
from __future__ import annotations

import numpy

from .nn_units import Forward, GradientDescentBase


def _sigm(x):
    return 1.0 / (1.0 + numpy.exp(-x))


class RecurrentBase(Forward):
    """Shared plumbing of RNN and LSTM forward units."""
    hide_from_registry = True
    GATES = 1
    PEER_WEIGHTS_OK = False

    def __init__(self, workflow, **kwargs):
        super(RecurrentBase, self).__init__(workflow, **kwargs)
        hidden = kwargs.get("hidden_size", kwargs.get("output_sample_shape", kwargs.get("output_samples_number")))
        if hidden is None:
            raise ValueError("hidden_size is required")
        if hasattr(hidden, "__iter__"):
            hidden = int(numpy.prod(tuple(hidden)))
        self.hidden_size = int(hidden)
        self.return_sequences = bool(kwargs.get("return_sequences", True))
        self.sequence_length = kwargs.get("sequence_length")
        self.exports = self.exports + ["hidden_size", "return_sequences", "sequence_length"]

    def init_unpickled(self):
        super(RecurrentBase, self).init_unpickled()
        self.np_cache_ = None
        self.xh_ = None

    # -- geometry ----------------------------------------------------------------------------------------------
    def geometry(self):
        """(batch, T, D) of the input."""
        shape = self.input.shape
        if len(shape) == 3 and self.sequence_length in (None, shape[1]):
            return int(shape[0]), int(shape[1]), int(shape[2])
        T = self.sequence_length
        if T is None:
            raise ValueError("%s: input of shape %s is not [batch, T, D]; pass sequence_length" % (self.name, shape))
        n = int(numpy.prod(shape[1:]))
        if n % int(T):
            raise ValueError("%s: sample size %d is not a multiple of sequence_length %d" % (self.name, n, T))
        return int(shape[0]), int(T), n // int(T)

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        B, T, D = self.geometry()
        H = self.hidden_size
        fresh = not self.weights
        self.create_weights(self.GATES * H, D + H)
        if fresh:
            self.init_bias()
        shape = (B, T, H) if self.return_sequences else (B, H)
        if not self.output or self.output.shape != shape:
            self.output.reset(numpy.zeros(shape, dtype=numpy.float32))
        self._backend_init_()
        return None

    def init_bias(self):
        pass

    # -- numpy oracle --------------------------------------------------------------------------------------
    def numpy_cell(self, z, state):
        raise NotImplementedError

    def numpy_run(self):
        B, T, D = self.geometry()
        H = self.hidden_size
        self.input.map_read()
        self.weights.map_read()
        x = self.input.mem.reshape(B, T, D).astype(numpy.float32)
        w = self.weights.mem
        b = 0.0
        if self.include_bias:
            self.bias.map_read()
            b = self.bias.mem
        xh = numpy.zeros((T, B, D + H), numpy.float32)
        xh[:, :, :D] = x.transpose(1, 0, 2)
        cache = {"xh": xh, "steps": []}
        state = None
        hs = numpy.zeros((T, B, H), numpy.float32)
        for t in range(T):
            z = xh[t].dot(w.T) + b
            h, state, saved = self.numpy_cell(z, state)
            cache["steps"].append(saved)
            hs[t] = h
            if t + 1 < T:
                xh[t + 1, :, D:] = h
        self.np_cache_ = cache
        self.output.map_invalidate()
        if self.return_sequences:
            self.output.mem[...] = hs.transpose(1, 0, 2)
        else:
            self.output.mem[...] = hs[-1]

    # -- device ---------------------------------------------------------------------------------------------------
    def b200_init(self):
        import torch
        from ..ops import kernels
        dev = self.device_
        B, T, D = self.geometry()
        H, G = self.hidden_size, self.GATES
        self.place_parameters()
        self.output.dev_dtype = dev.compute_dtype
        self.output.initialize(dev)
        tdev = dev.torch_device
        self.xh_ = kernels.pitched_empty(T * B, D + H, torch.bfloat16, tdev)          # zero: h_{-1} = 0
        self.z_ = kernels.pitched_empty(B, G * H, torch.float32, tdev, align=4)
        self.alloc_cell_state(B, T, H, tdev)

    def alloc_cell_state(self, B, T, H, tdev):
        raise NotImplementedError

    def device_cell(self, t, z, h_next, out):
        raise NotImplementedError

    def stage_input(self):
        """input [B, T, D] (any float dtype / pitch) -> the x columns of the time-major bf16 operand."""
        from ..ops import kernels
        B, T, D = self.geometry()
        x = self.input.devmem
        x2 = x if x.dim() == 2 else x.reshape(B, -1)
        for t in range(T):
            kernels.copy_2d(x2[:, t * D:(t + 1) * D], self.xh_[t * B:(t + 1) * B, :D])

    def b200_run(self):
        from ..ops import kernels
        self.input.unmap()
        B, T, D = self.geometry()
        H = self.hidden_size
        self.stage_input()
        out = self.output.devmem
        out2 = out if out.dim() == 2 else out.reshape(B, -1)
        bias = self.bias.devmem if self.include_bias else None
        for t in range(T):
            kernels.gemm(self.xh_[t * B:(t + 1) * B], self.weights_bf16_, self.z_, bias=bias)
            h_next = self.xh_[(t + 1) * B:(t + 2) * B, D:D + H] if t + 1 < T else None
            if self.return_sequences:
                o = out2[:, t * H:(t + 1) * H]
            else:
                o = out2 if t == T - 1 else None
            self.device_cell(t, self.z_, h_next, o)
        self.output.mark_device_written()


class RNN(RecurrentBase):
    """Elman recurrent layer, ``h_t = tanh([x_t | h_{t-1}] W^T + b)``."""
    hide_from_registry = False
    MAPPING = {"rnn"}
    GATES = 1
    __id__ = "0d2a3e57-7b0f-4c63-9b55-5a4e3d6f2c11"

    def numpy_cell(self, z, state):
        h = numpy.tanh(z)
        return h, h, h

    def alloc_cell_state(self, B, T, H, tdev):
        import torch
        self.hsave_ = torch.zeros(T, B, H, dtype=torch.float32, device=tdev)

    def device_cell(self, t, z, h_next, out):
        from ..ops import kernels
        kernels.rnn_cell_forward(z, self.hsave_[t], h_next, out)


class LSTM(RecurrentBase):
    """Long short-term memory layer (gate order i, f, g, o; ``forget_bias`` added to the f block at creation)."""
    hide_from_registry = False
    MAPPING = {"lstm"}
    GATES = 4
    __id__ = "6c1e8f0a-2d47-4b39-8a6e-9f3b7c5d1e24"

    def __init__(self, workflow, **kwargs):
        super(LSTM, self).__init__(workflow, **kwargs)
        self.forget_bias = float(kwargs.get("forget_bias", 1.0))

    def init_bias(self):
        if self.include_bias and self.bias and self.forget_bias:
            H = self.hidden_size
            self.bias.map_write()
            self.bias.mem[H:2 * H] += self.forget_bias

    def numpy_cell(self, z, state):
        H = self.hidden_size
        i, f, g, o = _sigm(z[:, :H]), _sigm(z[:, H:2 * H]), numpy.tanh(z[:, 2 * H:3 * H]), _sigm(z[:, 3 * H:])
        c_prev = state if state is not None else numpy.zeros_like(i)
        c = f * c_prev + i * g
        h = o * numpy.tanh(c)
        return h, c, (i, f, g, o, c_prev, c)

    def alloc_cell_state(self, B, T, H, tdev):
        import torch
        self.gates_ = torch.zeros(T, B, 4 * H, dtype=torch.float32, device=tdev)
        self.c_ = torch.zeros(T, B, H, dtype=torch.float32, device=tdev)

    def device_cell(self, t, z, h_next, out):
        from ..ops import kernels
        kernels.lstm_cell_forward(z, self.c_[t - 1] if t else None, self.c_[t], self.gates_[t], h_next, out)


class GDRecurrentBase(GradientDescentBase):
    """Back-propagation through time for RecurrentBase units (needs ``forward=`` the unit it differentiates)."""
    hide_from_registry = True
    CAN_FUSE_PREV_DERIVATIVE = False

    def __init__(self, workflow, **kwargs):
        super(GDRecurrentBase, self).__init__(workflow, **kwargs)
        self.demand("input", "output", "weights")

    def initialize(self, device=None, **kwargs):
        if self.err_output is None or not self.err_output or self.input is None or not self.input:
            return True
        if self.forward is None:
            raise ValueError("%s needs forward=<the recurrent unit> (its saved cell states)" % self.name)
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
    def numpy_cell_backward(self, dh, carry, saved):
        """-> (dz, carry')"""
        raise NotImplementedError

    def numpy_run(self):
        fwd = self.forward
        B, T, D = fwd.geometry()
        H = fwd.hidden_size
        cache = fwd.np_cache_
        self.err_output.map_read()
        eo = self.err_output.mem.astype(numpy.float32)
        if fwd.return_sequences:
            eo = eo.reshape(B, T, H).transpose(1, 0, 2)
        self.weights.map_write()
        w = self.weights.mem
        dz_all = numpy.zeros((T, B, fwd.GATES * H), numpy.float32)
        dx = numpy.zeros((T, B, D), numpy.float32)
        dh_rec = numpy.zeros((B, H), numpy.float32)
        carry = None
        for t in reversed(range(T)):
            dh = dh_rec.copy()
            if fwd.return_sequences:
                dh += eo[t]
            elif t == T - 1:
                dh += eo.reshape(B, H)
            dz, carry = self.numpy_cell_backward(dh, carry, cache["steps"][t])
            dz_all[t] = dz
            dxh = dz.dot(w)
            dx[t] = dxh[:, :D]
            dh_rec = dxh[:, D:]
        if self.need_err_input:
            self.err_input.map_invalidate()
            self.err_input.mem[...] = dx.transpose(1, 0, 2).reshape(self.err_input.shape)
        gw = dz_all.reshape(T * B, -1).T.dot(cache["xh"].reshape(T * B, -1))
        gw, = self.dp_reduce_host(gw)
        self.gradient_weights.reset(gw.astype(numpy.float32))
        if self.apply_gradient:
            self.numpy_update(w, gw, self.gradient_weights_with_moment.mem,
                              self.accumulated_gradient_weights.mem if self.accumulated_gradient_weights else None,
                              self.learning_rate, self.weights_decay, self.l1_vs_l2, self.gradient_moment)
        if self.bias is not None and self.bias:
            self.bias.map_write()
            gb = dz_all.reshape(T * B, -1).sum(axis=0)
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
        fwd = self.forward
        B, T, D = fwd.geometry()
        H, G = fwd.hidden_size, fwd.GATES
        self.place_state()
        if self.need_err_input:
            self.err_input.dev_dtype = dev.compute_dtype
            self.err_input.initialize(dev)
        tdev = dev.torch_device
        self.dz_ = kernels.pitched_empty(T * B, G * H, torch.bfloat16, tdev)
        self.dxh_ = kernels.pitched_empty(B, D + H, torch.float32, tdev, align=4)
        self.col_scratch_ = torch.empty(kernels.col_sum_scratch_elems(T * B, G * H), dtype=torch.float32, device=tdev)
        self.alloc_carry(B, H, tdev)

    def alloc_carry(self, B, H, tdev):
        pass

    def device_cell_backward(self, t, T, dh_ext, dh_rec, dz):
        raise NotImplementedError

    def b200_run(self):
        from ..ops import kernels
        self.err_output.unmap()
        fwd = self.forward
        B, T, D = fwd.geometry()
        H = fwd.hidden_size
        eo = self.err_output.devmem
        eo2 = eo if eo.dim() == 2 else eo.reshape(B, -1)
        ei2 = None
        if self.need_err_input:
            ei = self.err_input.devmem
            ei2 = ei if ei.dim() == 2 else ei.reshape(B, -1)
        w = fwd.weights_bf16_
        for t in reversed(range(T)):
            if fwd.return_sequences:
                dh_ext = eo2[:, t * H:(t + 1) * H]
            else:
                dh_ext = eo2 if t == T - 1 else None
            dh_rec = self.dxh_[:, D:D + H] if t + 1 < T else None
            dz = self.dz_[t * B:(t + 1) * B]
            self.device_cell_backward(t, T, dh_ext, dh_rec, dz)
            if t > 0 or self.need_err_input:
                kernels.gemm(dz, w, self.dxh_, b_mn=True)                   # [dx_t | dh_{t-1}] = dZ_t . W
                if self.need_err_input:
                    kernels.copy_2d(self.dxh_[:, :D], ei2[:, t * D:(t + 1) * D])
        if self.need_err_input:
            self.err_input.mark_device_written()
        gw_t, gb_t = self.grad_tensors()
        gw = gw_t.view(w.shape[0], -1)
        kernels.gemm(self.dz_, fwd.xh_, gw, a_mn=True, b_mn=True, accumulate=True, split_k=0)   # dW = dZ^T . xh
        if self.bias is not None and self.bias:
            self.device_bias_sum(self.dz_, gb_t)
        self.gradient_weights.mark_device_written()
        self.device_update()


class GDRNN(GDRecurrentBase):
    """Back-propagation through time for the simple recurrent layer."""
    hide_from_registry = False
    MAPPING = {"rnn"}

    def numpy_cell_backward(self, dh, carry, saved):
        return dh * (1.0 - saved * saved), None

    def device_cell_backward(self, t, T, dh_ext, dh_rec, dz):
        from ..ops import kernels
        kernels.rnn_cell_backward(dh_ext, dh_rec, self.forward.hsave_[t], dz)


class GDLSTM(GDRecurrentBase):
    """Back-propagation through time for the LSTM layer."""
    hide_from_registry = False
    MAPPING = {"lstm"}

    def numpy_cell_backward(self, dh, carry, saved):
        i, f, g, o, c_prev, c = saved
        tc = numpy.tanh(c)
        dc = (carry if carry is not None else 0.0) + dh * o * (1.0 - tc * tc)
        dz = numpy.concatenate([dc * g * i * (1 - i), dc * c_prev * f * (1 - f), dc * i * (1 - g * g),
                                dh * tc * o * (1 - o)], axis=1)
        return dz, dc * f

    def alloc_carry(self, B, H, tdev):
        import torch
        self.dc_ = torch.zeros(B, H, dtype=torch.float32, device=tdev)

    def device_cell_backward(self, t, T, dh_ext, dh_rec, dz):
        from ..ops import kernels
        fwd = self.forward
        kernels.lstm_cell_backward(dh_ext, dh_rec, self.dc_, t == T - 1, fwd.gates_[t],
                                   fwd.c_[t - 1] if t else None, fwd.c_[t], dz)
