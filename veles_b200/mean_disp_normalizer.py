"""MeanDispNormalizer unit: ``output = (input - mean) * rdisp`` on the device, uint8/float in, float out
(reference: ``veles/mean_disp_normalizer.py:50-138``; kernel ``cuda/mean_disp_normalizer.cu:11-19`` was
launched with one thread per block — here a vectorised grid-stride kernel, normally not needed at all
because the loader's gather kernel fuses the same arithmetic)."""
from __future__ import annotations

import numpy

from .accelerated_units import AcceleratedUnit
from .distributable import TriviallyDistributable
from .memory import Array


class MeanDispNormalizer(AcceleratedUnit, TriviallyDistributable):
    """output = (input - mean) * rdisp per feature, as a unit (device kernel `vb_mean_disp_normalize`)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "WORKER")
        super(MeanDispNormalizer, self).__init__(workflow, **kwargs)
        self.output = Array(shallow_pickle=True)
        self.global_size = None
        self.local_size = None
        self.demand("input", "mean", "rdisp")

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        for a in ("mean", "rdisp"):
            v = getattr(self, a)
            if not isinstance(v, Array):
                setattr(self, a, Array(numpy.asarray(v)))
        if self.mean.shape != self.rdisp.shape or self.mean.size != self.input.sample_size:
            raise ValueError("mean/rdisp shapes %s/%s do not match the sample size %d"
                             % (self.mean.shape, self.rdisp.shape, self.input.sample_size))
        if not self.output or self.output.shape != self.input.shape:
            self.output.reset(numpy.zeros(self.input.shape, dtype=numpy.float32))
        self._backend_init_()
        return None

    def numpy_run(self):
        self.input.map_read()
        self.mean.map_read()
        self.rdisp.map_read()
        self.output.map_invalidate()
        x = self.input.mem.reshape(self.input.shape[0], -1).astype(numpy.float32)
        out = (x - self.mean.mem.reshape(-1).astype(numpy.float32)) * self.rdisp.mem.reshape(-1)
        self.output.mem[...] = out.reshape(self.output.shape)

    def b200_init(self):
        import torch
        dev = self.device_
        self.init_vectors(self.input, self.output)
        self._mean_dev_ = torch.from_numpy(self.mean.mem.reshape(-1).astype(numpy.float32)).to(dev.torch_device)
        self._rdisp_dev_ = torch.from_numpy(self.rdisp.mem.reshape(-1).astype(numpy.float32)).to(dev.torch_device)

    def b200_run(self):
        from .ops import kernels
        self.input.unmap()
        out = self.output.devmem
        kernels.mean_disp_normalize(self.input.devmem, self._mean_dev_, self._rdisp_dev_,
                                    out.view(out.shape[0], -1))
        self.output.mark_device_written()
