"""Deconvolution (transposed convolution, ``deconv`` layer type) and its GD unit.

Forward of Deconv = backward-data of Conv: ``output = col2im(input · W)`` where ``input`` is [B,OH,OW,n_kernels]
and ``output`` [B,H,W,C]; weights have the Conv layout ``[n_kernels, ky*kx*C]`` so a Deconv can share ("tie")
weights with the Conv it inverts (autoencoders).  Backward: ``err_input = im2col(err_output) · Wᵀ``,
``dW = inputᵀ · im2col(err_output)``.  Device path reuses the tcgen05 GEMMs + im2col/col2im kernels."""
from __future__ import annotations

import numpy

from ..memory import Array
from .conv import numpy_im2col, numpy_col2im, conv_output_size
from .nn_units import Forward, GradientDescentBase


class Deconv(Forward):
    """Transposed convolution sharing (or owning) the weights of a Conv layer: convolutional auto-encoders."""
    hide_from_registry = False
    MAPPING = {"deconv"}
    ACTIVATION = 0

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("include_bias", False)
        super(Deconv, self).__init__(workflow, **kwargs)
        self.n_kernels = int(kwargs["n_kernels"])
        self.kx, self.ky = int(kwargs["kx"]), int(kwargs["ky"])
        self.padding = tuple(kwargs.get("padding", (0, 0, 0, 0)))
        self.sliding = tuple(kwargs.get("sliding", (1, 1)))
        self.n_channels = kwargs.get("n_channels")          # C of the output; or give output_shape_source
        self.output_shape_source = kwargs.get("output_shape_source")  # Array whose shape the output must match
        self.grouping = 1

    def init_unpickled(self):
        super(Deconv, self).init_unpickled()
        self.cols_ = None

    def out_shape4(self):
        B, OH, OW, nk = self.input.shape
        if nk != self.n_kernels:
            raise ValueError("Deconv input has %d channels, n_kernels is %d" % (nk, self.n_kernels))
        if self.output_shape_source is not None and self.output_shape_source:
            s = self.output_shape_source.shape
            return (B,) + tuple(s[1:]) if len(s) == 4 else (B,) + tuple(s[1:]) + (1,)
        pl, pt, pr, pb = self.padding
        sx, sy = self.sliding
        H = (OH - 1) * sy + self.ky - pt - pb
        W = (OW - 1) * sx + self.kx - pl - pr
        if self.n_channels is None:
            raise ValueError("Deconv needs n_channels or output_shape_source")
        return (B, H, W, int(self.n_channels))

    @property
    def geometry(self):
        B, H, W, C = self.output.shape
        _, OH, OW, _ = self.input.shape
        pl, pt, pr, pb = self.padding
        sx, sy = self.sliding
        return dict(OH=OH, OW=OW, ky=self.ky, kx=self.kx, sy=sy, sx=sx, pt=pt, pl=pl, groups=1)

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        if self.output_shape_source is not None and not self.output_shape_source:
            return True
        self.device = device
        shape = self.out_shape4()
        B, H, W, C = shape
        pl, pt, pr, pb = self.padding
        sx, sy = self.sliding
        if conv_output_size(H, self.ky, pt, pb, sy) != self.input.shape[1] or \
                conv_output_size(W, self.kx, pl, pr, sx) != self.input.shape[2]:
            raise ValueError("Deconv geometry mismatch: output %s does not convolve back to input %s"
                             % (shape, self.input.shape))
        self.create_weights(self.n_kernels, self.ky * self.kx * C)
        self.alloc_output(shape)
        self._backend_init_()
        return None

    def numpy_run(self):
        self.input.map_read()
        self.weights.map_read()
        x = self.input.mem.reshape(-1, self.n_kernels).astype(numpy.float32)
        cols = x.dot(self.weights.mem)
        pl, pt, pr, pb = self.padding
        sx, sy = self.sliding
        out = numpy_col2im(cols, self.output.shape, self.ky, self.kx, sy, sx, pl, pt, pr, pb)
        self.output.map_invalidate()
        self.output.mem[...] = out

    def b200_init(self):
        import torch
        dev = self.device_
        self.place_parameters()
        self.output.dev_dtype = dev.compute_dtype
        self.output.initialize(dev)
        kc = self.weights_bf16_.shape[1]
        rows = self.input.size // self.n_kernels
        self.cols_ = torch.zeros(rows, (kc + 7) // 8 * 8, dtype=torch.bfloat16, device=dev.torch_device)

    def b200_run(self):
        from ..ops import kernels
        x = self.input.devmem.reshape(-1, self.n_kernels)
        kc = self.weights_bf16_.shape[1]
        kernels.gemm(x, self.weights_bf16_, self.cols_[:, :kc], b_mn=True)
        kernels.col2im(self.cols_, self.output.devmem.reshape(self.output.shape), self.geometry)
        self.output.mark_device_written()


class GDDeconv(GradientDescentBase):
    """Backward pass of Deconv."""
    hide_from_registry = False
    MAPPING = {"deconv"}

    def __init__(self, workflow, **kwargs):
        super(GDDeconv, self).__init__(workflow, **kwargs)
        self.demand("input", "output", "weights")

    def initialize(self, device=None, **kwargs):
        if self.err_output is None or not self.err_output or self.input is None or not self.input:
            return True
        self.device = device
        self.alloc_err_input()
        self._ensure_state(self.weights, self.gradient_weights_with_moment, self.accumulated_gradient_weights)
        self._backend_init_()
        return None

    def numpy_run(self):
        f = self.forward
        self.err_output.map_read()
        self.input.map_read()
        self.weights.map_write()
        eo = self.err_output.mem.reshape(f.output.shape).astype(numpy.float32)
        pl, pt, pr, pb = f.padding
        sx, sy = f.sliding
        ecols, _, _ = numpy_im2col(eo, f.ky, f.kx, sy, sx, pl, pt, pr, pb)
        x = self.input.mem.reshape(-1, f.n_kernels).astype(numpy.float32)
        w = self.weights.mem
        if self.need_err_input:
            self.err_input.map_invalidate()
            self.err_input.mem[...] = ecols.dot(w.T).reshape(self.err_input.shape)
        gw = x.T.dot(ecols)
        self.gradient_weights.reset(gw.astype(numpy.float32))
        if self.apply_gradient:
            self.numpy_update(w, gw, self.gradient_weights_with_moment.mem,
                              self.accumulated_gradient_weights.mem if self.accumulated_gradient_weights else None,
                              self.learning_rate, self.weights_decay, self.l1_vs_l2, self.gradient_moment)

    def b200_init(self):
        dev = self.device_
        self.place_state()
        if self.need_err_input:
            self.err_input.dev_dtype = dev.compute_dtype
            self.err_input._pitched = False
            self.err_input.initialize(dev)

    def b200_run(self):
        from ..ops import kernels
        f = self.forward
        eo = self.err_output.devmem.reshape(f.output.shape)
        cols = f.cols_
        kc = f.weights_bf16_.shape[1]
        kernels.im2col(eo.contiguous(), cols, f.geometry)
        x = self.input.devmem.reshape(-1, f.n_kernels)
        if self.need_err_input:
            kernels.gemm(cols[:, :kc], f.weights_bf16_, self.err_input.devmem.reshape(-1, f.n_kernels))
            self.err_input.mark_device_written()
        gw_t, _ = self.grad_tensors()
        kernels.gemm(x, cols[:, :kc], gw_t.view(f.n_kernels, kc), a_mn=True, b_mn=True, accumulate=True, split_k=0)
        self.gradient_weights.mark_device_written()
        self.device_update()
