"""Convolutional layers (NHWC): ``conv``, ``conv_tanh``, ``conv_relu``, ``conv_str``, ``conv_sigmoid``.

kwargs (``docs/source/manualrst_veles_workflow_parameters.rst:518-541``): ``n_kernels, kx, ky,
padding (left, top, right, bottom), sliding (x, y), grouping, unpack_size, weights_filling,
weights_stddev, bias_filling, bias_stddev, include_bias, weights_transposed``.
Pinned layout: input ``[B, H, W, C]``, weights ``[n_kernels, ky*kx*C/grouping]`` with column
order (tap row, tap col, channel), output ``[B, OH, OW, n_kernels]``; with ``grouping = G``
kernels ``g*n/G .. (g+1)*n/G`` see input channels ``g*C/G .. (g+1)*C/G``.

Device path: an implicit GEMM.  The tcgen05 kernel's TMA producer walks the NHWC input in im2col
mode (``cp.async.bulk.tensor.4d...im2col``: one box = 128 output pixels x 64 channels of one filter
tap, out-of-image taps read as zeros), so the column matrix Znicz "unpacked" in ``unpack_size``
chunks never exists; the bias+activation epilogue writes straight into the NHWC output (a column
slice per group).  Layers whose per-group channel count is not a multiple of 8 (the 3-channel first
layer) use one vectorised im2col into a resident bf16 column matrix and the same GEMM.
"""
from __future__ import annotations

import numpy

from ..config import root, get
from ..memory import Array
from .nn_units import Forward, act_forward_numpy


def conv_output_size(size, k, pad_lo, pad_hi, stride):
    return (size + pad_lo + pad_hi - k) // stride + 1


def numpy_im2col(x, ky, kx, sy, sx, pl, pt, pr, pb, groups=1):
    """x [B,H,W,C] -> cols [B*OH*OW, G*ky*kx*Cg] in (group, tap, channel) order."""
    B, H, W, C = x.shape
    xp = numpy.pad(x, ((0, 0), (pt, pb), (pl, pr), (0, 0)))
    OH = (H + pt + pb - ky) // sy + 1
    OW = (W + pl + pr - kx) // sx + 1
    s = xp.strides
    win = numpy.lib.stride_tricks.as_strided(
        xp, shape=(B, OH, OW, ky, kx, C), strides=(s[0], s[1] * sy, s[2] * sx, s[1], s[2], s[3]), writeable=False)
    Cg = C // groups
    win = win.reshape(B * OH * OW, ky * kx, groups, Cg).transpose(0, 2, 1, 3)
    return numpy.ascontiguousarray(win).reshape(B * OH * OW, groups * ky * kx * Cg), OH, OW


def numpy_col2im(dcols, shape, ky, kx, sy, sx, pl, pt, pr, pb, groups=1):
    """Adjoint of numpy_im2col."""
    B, H, W, C = shape
    OH = (H + pt + pb - ky) // sy + 1
    OW = (W + pl + pr - kx) // sx + 1
    Cg = C // groups
    d = dcols.reshape(B, OH, OW, groups, ky, kx, Cg).transpose(0, 1, 2, 4, 5, 3, 6).reshape(B, OH, OW, ky, kx, C)
    out = numpy.zeros((B, H + pt + pb, W + pl + pr, C), dtype=dcols.dtype)
    for i in range(ky):
        for j in range(kx):
            out[:, i:i + sy * OH:sy, j:j + sx * OW:sx, :] += d[:, :, :, i, j, :]
    return out[:, pt:pt + H, pl:pl + W, :]


class Conv(Forward):
    """Convolutional layer, linear output: NHWC input, `n_kernels` filters of `ky` x `kx`, `sliding`,
    `padding`, `grouping`; implicit-GEMM tcgen05 kernels on the device."""
    hide_from_registry = False
    MAPPING = {"conv"}
    ACTIVATION = 0
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0001"

    def __init__(self, workflow, **kwargs):
        super(Conv, self).__init__(workflow, **kwargs)
        try:
            self.n_kernels = int(kwargs["n_kernels"])
            self.kx = int(kwargs["kx"])
            self.ky = int(kwargs["ky"])
        except KeyError:
            raise KeyError("n_kernels, kx and ky are required parameters of %s" % type(self).__name__)
        self.padding = tuple(kwargs.get("padding", (0, 0, 0, 0)))  # left, top, right, bottom
        self.sliding = tuple(kwargs.get("sliding", (1, 1)))        # x, y
        self.grouping = int(kwargs.get("grouping", 1))
        self.unpack_size = kwargs.get("unpack_size", 16)           # accepted for config compatibility
        if len(self.padding) != 4 or len(self.sliding) != 2:
            raise ValueError("padding must be (l, t, r, b) and sliding (x, y)")
        self.exports += ["n_kernels", "kx", "ky", "padding", "sliding", "grouping"]

    def init_unpickled(self):
        super(Conv, self).init_unpickled()
        self.cols_ = None

    @property
    def in_shape4(self):
        s = self.input.shape
        if len(s) == 4:
            return s
        if len(s) == 3:
            return s + (1,)
        raise ValueError("Conv input must be [B,H,W,C] or [B,H,W], got %s" % (s,))

    @property
    def geometry(self):
        B, H, W, C = self.in_shape4
        pl, pt, pr, pb = self.padding
        sx, sy = self.sliding
        return dict(OH=conv_output_size(H, self.ky, pt, pb, sy), OW=conv_output_size(W, self.kx, pl, pr, sx),
                    ky=self.ky, kx=self.kx, sy=sy, sx=sx, pt=pt, pl=pl, groups=self.grouping, H=H, W=W)

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        B, H, W, C = self.in_shape4
        if C % self.grouping or self.n_kernels % self.grouping:
            raise ValueError("channels (%d) and n_kernels (%d) must be divisible by grouping (%d)"
                             % (C, self.n_kernels, self.grouping))
        g = self.geometry
        if g["OH"] < 1 or g["OW"] < 1:
            raise ValueError("Convolution kernel does not fit the padded input")
        self.create_weights(self.n_kernels, self.ky * self.kx * C // self.grouping)
        self.alloc_output((B, g["OH"], g["OW"], self.n_kernels))
        self._backend_init_()
        return None

    def default_weights_stddev(self, fan_in, fan_out):
        return float(min(numpy.sqrt(6.0 / (fan_in + fan_out * self.kx * self.ky // max(self.sliding[0] * self.sliding[1], 1))), 0.5))

    # -- numpy oracle -----------------------------------------------------------------------------------------
    def numpy_run(self):
        self.input.map_read()
        self.weights.map_read()
        x = self.input.mem.reshape(self.in_shape4).astype(numpy.float32)
        pl, pt, pr, pb = self.padding
        sx, sy = self.sliding
        G = self.grouping
        cols, OH, OW = numpy_im2col(x, self.ky, self.kx, sy, sx, pl, pt, pr, pb, G)
        w = self.weights.mem if not self.weights_transposed else self.weights.mem.T
        nkg = self.n_kernels // G
        kg = cols.shape[1] // G
        out = numpy.empty((cols.shape[0], self.n_kernels), dtype=numpy.float32)
        for gi in range(G):
            out[:, gi * nkg:(gi + 1) * nkg] = cols[:, gi * kg:(gi + 1) * kg].dot(w[gi * nkg:(gi + 1) * nkg].T)
        if self.include_bias:
            self.bias.map_read()
            out += self.bias.mem
        self.output.map_invalidate()
        self.output.mem[...] = act_forward_numpy(self.ACTIVATION, out).reshape(self.output.shape)

    # -- device ---------------------------------------------------------------------------------------------------------
    def b200_init(self):
        import torch
        dev = self.device_
        self.place_parameters()
        self.output.dev_dtype = dev.compute_dtype
        self.output.initialize(dev)
        B, H, W, C = self.in_shape4
        g = self.geometry
        kk = self.ky * self.kx * (C // self.grouping)
        if self.grouping > 1 and (kk % 8 or (self.n_kernels // self.grouping) % 8):
            raise ValueError("grouped convolution on the device needs ky*kx*C/G and n_kernels/G to be multiples of 8")
        # Implicit GEMM (TMA im2col-mode loads, no column matrix) whenever the per-group channel count allows
        # 16-byte channel vectors; otherwise (e.g. the 3-channel first layer) an explicit bf16 column matrix.
        Cg, nkg = C // self.grouping, self.n_kernels // self.grouping
        self.implicit_ = (Cg % 8 == 0 and nkg % 8 == 0 and not getattr(self, "force_explicit_im2col", False))
        self.cols_ld_ = (self.grouping * kk + 7) // 8 * 8
        self.cols_ = None
        # Strided few-channel layers (11x11/4 on RGB): space-to-depth by the stride makes them unit-stride
        # convolutions with ceil(k/s)^2 taps over sy*sx*C channels, which the implicit GEMM can walk.
        sx, sy = self.sliding
        self.s2d_ = None
        if (not self.implicit_ and self.grouping == 1 and sx * sy > 1 and not any(self.padding)
                and (sx * sy * C) % 8 == 0 and self.n_kernels % 8 == 0
                and not getattr(self, "force_explicit_im2col", False)):
            self._init_space_to_depth(B, H, W, C, g)
        if not self.implicit_ and self.s2d_ is None:
            self.ensure_cols()
        self._x_stage4_ = None
        if self.input.devmem.dtype != torch.bfloat16:
            self._x_stage4_ = torch.zeros(self.in_shape4, dtype=torch.bfloat16, device=dev.torch_device)

    def _init_space_to_depth(self, B, H, W, C, g):
        import torch
        dev = self.device_.torch_device
        sx, sy = self.sliding
        ky2, kx2 = (self.ky + sy - 1) // sy, (self.kx + sx - 1) // sx
        H2, W2, C2 = (H + sy - 1) // sy, (W + sx - 1) // sx, sy * sx * C
        # column maps between the Znicz weight layout [(r, s, c)] and the space-to-depth one [(by, bx, dy, dx, c)]
        to_s2d = numpy.full(ky2 * kx2 * C2, -1, dtype=numpy.int32)
        from_s2d = numpy.zeros(self.ky * self.kx * C, dtype=numpy.int32)
        for by in range(ky2):
            for bx in range(kx2):
                for dy in range(sy):
                    for dx in range(sx):
                        r, s_ = by * sy + dy, bx * sx + dx
                        if r >= self.ky or s_ >= self.kx:
                            continue
                        for c in range(C):
                            j2 = ((by * kx2 + bx) * sy * sx + dy * sx + dx) * C + c
                            j = (r * self.kx + s_) * C + c
                            to_s2d[j2] = j
                            from_s2d[j] = j2
        self.s2d_ = dict(
            geom=dict(OH=g["OH"], OW=g["OW"], ky=ky2, kx=kx2, sy=1, sx=1, pt=0, pl=0, groups=1, H=H2, W=W2),
            x=torch.zeros(B, H2, W2, C2, dtype=torch.bfloat16, device=dev),
            w=torch.zeros(self.n_kernels, ky2 * kx2 * C2, dtype=torch.bfloat16, device=dev),
            gw=torch.zeros(self.n_kernels, ky2 * kx2 * C2, dtype=torch.float32, device=dev),
            to_s2d=torch.from_numpy(to_s2d).to(dev), from_s2d=torch.from_numpy(from_s2d).to(dev))
        self._offer_input_sink(B, H, W, C)

    def _offer_input_sink(self, B, H, W, C):
        """When this layer is the ONLY reader of the loader's minibatch, the loader gathers straight into the
        space-to-depth layout (``vb_gather_s2d``): the NHWC minibatch is neither written nor re-read.  Any other unit
        that holds the same Array (a plotter, an image saver, an MSE evaluator...) keeps the two-kernel path."""
        if not get(root.common.engine.fuse_input_gather, True):
            return
        wf = self.workflow
        loader = getattr(wf, "loader", None)
        if loader is None or getattr(loader, "minibatch_data", None) is not self.input:
            return
        if self.input.devmem is None or tuple(self.input.shape[1:]) != (H, W, C):
            return
        for u in list(getattr(wf, "_units", [])):
            if u is loader or u is self or getattr(u, "forward", None) is self:
                continue       # (this layer's GD unit takes its weight gradient from the rearranged input too)
            if any(v is self.input for v in u.__dict__.values()):
                return
        sx, sy = self.sliding
        loader.minibatch_sink_ = dict(xs=self.s2d_["x"], H=H, W=W, C=C, sy=sy, sx=sx, fed=False, owner=self.name)
        self.s2d_["sink"] = loader.minibatch_sink_

    def ensure_cols(self):
        """The [B*OH*OW, G*ky*kx*Cg] bf16 column matrix (explicit path and strided err_input scratch)."""
        import torch
        if self.cols_ is None:
            g = self.geometry
            self.cols_ = torch.zeros(self.in_shape4[0] * g["OH"] * g["OW"], self.cols_ld_, dtype=torch.bfloat16,
                                     device=self.device_.torch_device)
        return self.cols_

    def device_input4(self):
        from ..ops import kernels
        x = self.input.devmem
        B = self.in_shape4[0]
        if self._x_stage4_ is not None:
            kernels.copy_2d(x.reshape(B, -1), self._x_stage4_.view(B, -1))
            return self._x_stage4_
        return x.reshape(self.in_shape4) if tuple(x.shape) != tuple(self.in_shape4) else x

    def b200_run(self):
        from ..ops import kernels
        self.input.unmap()
        x = self.device_input4()
        g = self.geometry
        G = self.grouping
        nkg = self.n_kernels // G
        kg = self.weights_bf16_.shape[1]
        out2 = self.output.devmem.view(-1, self.n_kernels)
        bias = self.bias.devmem if self.include_bias else None
        if self.s2d_ is not None:
            sd = self.s2d_
            if not (sd.get("sink") is not None and sd["sink"]["fed"]):
                kernels.space_to_depth(x, sd["x"], g["sy"], g["sx"])
            kernels.gather_cols(self.weights_bf16_, sd["w"], sd["to_s2d"])   # refreshed bf16 weights, s2d order
            if not (kernels.conv_halo_preferred(sd["x"].shape[3], sd["geom"]) and
                    kernels.conv_forward_halo(sd["x"], sd["w"], out2, sd["geom"], bias=bias, act=self.ACTIVATION)):
                kernels.conv_forward(sd["x"], sd["w"], out2, sd["geom"], bias=bias, act=self.ACTIVATION)
        elif self.implicit_ and kernels.grouped_conv_ok(x.shape[3], self.n_kernels, G):
            # every group in ONE launch (tile -> group): fewer, fuller waves than one launch per group
            kernels.conv_forward_grouped(x, self.weights_bf16_, out2, g, G, bias=bias, act=self.ACTIVATION)
        elif self.implicit_:
            Cg = x.shape[3] // G
            halo = kernels.conv_halo_preferred(Cg, g)
            for gi in range(G):
                args = (x[..., gi * Cg:(gi + 1) * Cg] if G > 1 else x, self.weights_bf16_[gi * nkg:(gi + 1) * nkg],
                        out2[:, gi * nkg:(gi + 1) * nkg] if G > 1 else out2, g)
                kw = dict(bias=bias[gi * nkg:(gi + 1) * nkg] if bias is not None else None, act=self.ACTIVATION)
                if not (halo and kernels.conv_forward_halo(*args, **kw)):
                    kernels.conv_forward(*args, **kw)
        else:
            kernels.im2col(x, self.cols_, g)
            for gi in range(G):
                kernels.gemm(self.cols_[:, gi * kg:(gi + 1) * kg], self.weights_bf16_[gi * nkg:(gi + 1) * nkg],
                             out2[:, gi * nkg:(gi + 1) * nkg] if G > 1 else out2,
                             bias=bias[gi * nkg:(gi + 1) * nkg] if bias is not None else None, act=self.ACTIVATION)
        self.output.mark_device_written()


class ConvTanh(Conv):
    """Convolutional layer with scaled tanh activation."""
    MAPPING = {"conv_tanh"}
    ACTIVATION = 1
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0002"


class ConvRELU(Conv):
    """Convolutional layer with softplus activation."""
    MAPPING = {"conv_relu"}
    ACTIVATION = 3
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0004"


class ConvStrictRELU(Conv):
    """Convolutional layer with y = max(0, .)."""
    MAPPING = {"conv_str"}
    ACTIVATION = 4
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0005"


class ConvSigmoid(Conv):
    """Convolutional layer with logistic activation."""
    MAPPING = {"conv_sigmoid"}
    ACTIVATION = 2
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0003"
