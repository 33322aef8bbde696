"""Pooling layers (NHWC): ``max_pooling``, ``maxabs_pooling``, ``avg_pooling``,
``stochastic_pooling``, ``stochastic_abs_pooling`` and the pool-depool pairs
(``docs/source/manualrst_veles_workflow_parameters.rst:471-502``; kwargs ``kx, ky, sliding``).
Pinned semantics: windows start at multiples of ``sliding`` and are clipped at the right /
bottom border (``OH = ceil((H - ky) / sy) + 1``); max-like kinds record ``input_offset`` =
flat index of the chosen input element, which makes the backward pass an exact routing.
"""
from __future__ import annotations

import numpy

from .. import prng
from ..memory import Array
from .nn_units import ForwardBase


def pool_output_size(size, k, stride):
    if size <= k:
        return 1
    return -(-(size - k) // stride) + 1


class Pooling(ForwardBase):
    hide_from_registry = True
    KIND = "max"

    def __init__(self, workflow, **kwargs):
        super(Pooling, self).__init__(workflow, **kwargs)
        self.kx = int(kwargs.get("kx", 2))
        self.ky = int(kwargs.get("ky", 2))
        self.sliding = tuple(kwargs.get("sliding", (self.kx, self.ky)))
        self.input_offset = Array(shallow_pickle=True)
        self.exports = ["kx", "ky", "sliding"]

    @property
    def in_shape4(self):
        s = self.input.shape
        return s if len(s) == 4 else s + (1,)

    @property
    def out_shape4(self):
        B, H, W, C = self.in_shape4
        return (B, pool_output_size(H, self.ky, self.sliding[1]), pool_output_size(W, self.kx, self.sliding[0]), C)

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        self.alloc_output(self.out_shape4)
        if self.KIND != "avg" and (not self.input_offset or self.input_offset.shape != self.out_shape4):
            self.input_offset.reset(numpy.zeros(self.out_shape4, dtype=numpy.int32))
        self._backend_init_()
        return None

    def package_export(self):
        return {"kx": self.kx, "ky": self.ky, "sliding": list(self.sliding), "kind": self.KIND}

    # -- numpy oracle -------------------------------------------------------------------------------------------
    def _windows(self):
        B, H, W, C = self.in_shape4
        _, OH, OW, _ = self.out_shape4
        sx, sy = self.sliding
        for oh in range(OH):
            for ow in range(OW):
                h0, w0 = oh * sy, ow * sx
                yield oh, ow, h0, min(h0 + self.ky, H), w0, min(w0 + self.kx, W)

    def numpy_run(self):
        self.input.map_read()
        x = self.input.mem.reshape(self.in_shape4).astype(numpy.float32)
        B, H, W, C = x.shape
        self.output.map_invalidate()
        out = self.output.mem.reshape(self.out_shape4)
        offs = None
        if self.KIND != "avg":
            self.input_offset.map_invalidate()
            offs = self.input_offset.mem
        rnd = self._random_u16() if self.KIND.startswith("stochastic") else None
        bidx = numpy.arange(B)[:, None]
        cidx = numpy.arange(C)[None, :]
        for oh, ow, h0, h1, w0, w1 in self._windows():
            win = x[:, h0:h1, w0:w1, :].reshape(B, -1, C)
            ww = w1 - w0
            if self.KIND == "avg":
                out[:, oh, ow, :] = win.mean(axis=1)
                continue
            if self.KIND in ("max", "maxabs"):
                key = numpy.abs(win) if self.KIND == "maxabs" else win
                am = key.argmax(axis=1)
            else:
                key = numpy.abs(win) if self.KIND == "stochastic_abs" else numpy.maximum(win, 0)
                cs = numpy.cumsum(key, axis=1)
                total = cs[:, -1, :]
                pos = total * (rnd[:, oh, ow, :].astype(numpy.float32) / 65536.0)
                am = (cs < pos[:, None, :]).sum(axis=1)
                am = numpy.minimum(am, win.shape[1] - 1)
                am[total <= 0] = 0
            out[:, oh, ow, :] = win[bidx, am, cidx]
            hh = h0 + am // ww
            wi = w0 + am % ww
            offs[:, oh, ow, :] = ((bidx * H + hh) * W + wi) * C + cidx

    def _random_u16(self):
        return self.rand.randint(0, 65536, size=self.out_shape4).astype(numpy.uint16)

    # -- device -----------------------------------------------------------------------------------------------------
    def b200_init(self):
        dev = self.device_
        self.output.dev_dtype = self.input.devmem.dtype if self.input.devmem is not None else dev.compute_dtype
        self.output.initialize(dev)
        if self.KIND != "avg":
            self.input_offset.initialize(dev)
        # Hot path of max / maxabs pooling: one byte per output (window-local argmax) is all the backward pass
        # needs; the Znicz-visible int32 ``input_offset`` is written only for units that asked for it
        # (``need_input_offset``) and otherwise expanded lazily when the host maps it.
        self.offs8_ = None
        self.offsets_stale_ = False
        import torch
        if (self.KIND in ("max", "maxabs") and self.output.devmem.dtype == torch.bfloat16
                and self.in_shape4[3] % 8 == 0 and self.kx * self.ky <= 255):
            self.offs8_ = torch.zeros(self.out_shape4, dtype=torch.uint8, device=dev.torch_device)
            self.input_offset.lazy_device_ = self.ensure_input_offset

    def ensure_input_offset(self):
        """Bring the int32 ``input_offset`` up to date on the device (no-op unless the compact path skipped it)."""
        if getattr(self, "offs8_", None) is None or not self.offsets_stale_:
            return
        from ..ops import kernels
        self.offsets_stale_ = False
        kernels.pool_expand_offsets(self.offs8_, self.input_offset.devmem_, self.in_shape4, self.ky, self.kx,
                                    self.sliding[1], self.sliding[0])
        self.input_offset.mark_device_written()

    def b200_run(self):
        import torch
        from ..ops import kernels
        self.input.unmap()
        x = self.input.devmem.reshape(self.in_shape4)
        y = self.output.devmem.reshape(self.out_shape4)
        rnd = None
        if self.KIND.startswith("stochastic"):
            rnd = torch.from_numpy(self._random_u16().view(numpy.int16)).to(x.device)
        if self.offs8_ is not None:
            want_int = bool(getattr(self, "need_input_offset", False))
            if kernels.pool_forward_local(kernels.POOL_KINDS[self.KIND], x, y,
                                          self.input_offset.devmem_ if want_int else None, self.offs8_,
                                          self.ky, self.kx, self.sliding[1], self.sliding[0]):
                self.output.mark_device_written()
                self.offsets_stale_ = not want_int
                if want_int:
                    self.input_offset.mark_device_written()
                return
            self.offs8_ = None   # shape not supported by the compact path: use the int32 one from now on
        kernels.pool_forward(kernels.POOL_KINDS[self.KIND], x, y,
                             self.input_offset.devmem if self.KIND != "avg" else None,
                             self.ky, self.kx, self.sliding[1], self.sliding[0], rnd)
        self.output.mark_device_written()
        if self.KIND != "avg":
            self.input_offset.mark_device_written()


class MaxPooling(Pooling):
    """Max pooling; records the chosen position of every window for the backward pass."""
    hide_from_registry = False
    MAPPING = {"max_pooling"}
    KIND = "max"
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0011"


class MaxAbsPooling(Pooling):
    """Pooling that keeps the element of largest magnitude (with its sign)."""
    hide_from_registry = False
    MAPPING = {"maxabs_pooling"}
    KIND = "maxabs"
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0012"


class AvgPooling(Pooling):
    """Average pooling over `ky` x `kx` windows with `sliding`."""
    hide_from_registry = False
    MAPPING = {"avg_pooling"}
    KIND = "avg"
    __id__ = "0e4d2e9a-6a0b-4c1b-9a5f-1c4b9c7e0013"


class StochasticPoolingBase(Pooling):
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(StochasticPoolingBase, self).__init__(workflow, **kwargs)
        self.rand = kwargs.get("rand", prng.get())


class StochasticPooling(StochasticPoolingBase):
    """Pooling that samples an element of the window with probability proportional to its value."""
    hide_from_registry = False
    MAPPING = {"stochastic_pooling"}
    KIND = "stochastic"


class StochasticAbsPooling(StochasticPoolingBase):
    """Stochastic pooling on magnitudes."""
    hide_from_registry = False
    MAPPING = {"stochastic_abs_pooling"}
    KIND = "stochastic_abs"


class StochasticPoolingDepooling(StochasticPoolingBase):
    """Pool then immediately depool: output has the INPUT shape, zero except the chosen elements."""
    hide_from_registry = False
    MAPPING = {"stochastic_pool_depool"}
    KIND = "stochastic"

    def initialize(self, device=None, **kwargs):
        res = super(StochasticPoolingDepooling, self).initialize(device=device, **kwargs)
        if res:
            return res
        self._pooled = Array(shallow_pickle=True)
        self._pooled.reset(numpy.zeros(self.out_shape4, dtype=numpy.float32))
        self.output.reset(numpy.zeros(self.in_shape4, dtype=numpy.float32))
        if self.on_device:
            self.output.dev_dtype = self.input.devmem.dtype
            self.output.initialize(self.device_)
            self._pooled.dev_dtype = self.input.devmem.dtype
            self._pooled.initialize(self.device_)
        return None

    def numpy_run(self):
        real_out = self.output
        self.output = self._pooled
        try:
            super(StochasticPoolingDepooling, self).numpy_run()
        finally:
            self.output = real_out
        self.output.map_invalidate()
        self.output.mem[...] = 0
        self.input_offset.map_read()
        self.output.mem.reshape(-1)[self.input_offset.mem.reshape(-1)] = self._pooled.mem.reshape(-1)

    def b200_run(self):
        real_out = self.output
        self.output = self._pooled
        try:
            super(StochasticPoolingDepooling, self).b200_run()
        finally:
            self.output = real_out
        out = self.output.devmem
        out.zero_()
        out.view(-1)[self.input_offset.devmem.view(-1).long()] = self._pooled.devmem.view(-1)
        self.output.mark_device_written()


class StochasticAbsPoolingDepooling(StochasticPoolingDepooling):
    """Stochastic-abs pooling immediately undone in place (a sparsifying regulariser)."""
    MAPPING = {"stochastic_abs_pool_depool"}
    KIND = "stochastic_abs"
