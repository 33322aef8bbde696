"""InputJoiner: concatenates several inputs sample-wise into one output
(reference: ``veles/input_joiner.py:49-212``; kernel ``cuda/join.jcu`` was a Jinja2 template with one
pointer argument per input and ONE thread per sample; here each input is a strided 2-D copy)."""
from __future__ import annotations

import numpy

from .accelerated_units import AcceleratedUnit
from .distributable import TriviallyDistributable
from .memory import Array


class InputJoiner(AcceleratedUnit, TriviallyDistributable):
    """Concatenates the samples of several input Arrays along the feature axis into one output (column blocks
    copied by `vb_copy_2d` on the device)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "WORKER")
        super(InputJoiner, self).__init__(workflow, **kwargs)
        self.inputs = list(kwargs.get("inputs", []))
        self.num_inputs = kwargs.get("num_inputs", len(self.inputs))
        self.output = Array(shallow_pickle=True)
        self.offsets = []
        self.lengths = []
        self.minibatch_size = None

    def link_inputs(self, other, *attrs):
        """Append ``other.<attr>`` Arrays to the list of inputs (mutable => shared by reference)."""
        for attr in attrs:
            self.inputs.append(getattr(other, attr))
        self.num_inputs = len(self.inputs)
        return self

    def initialize(self, device=None, **kwargs):
        if not self.inputs or any(i is None or not i for i in self.inputs):
            return True
        self.device = device
        batch = self.inputs[0].shape[0]
        if any(i.shape[0] != batch for i in self.inputs):
            raise ValueError("All inputs must have the same number of samples")
        self.lengths = [i.sample_size for i in self.inputs]
        self.offsets = [int(sum(self.lengths[:k])) for k in range(len(self.lengths))]
        total = int(sum(self.lengths))
        if not self.output or self.output.shape != (batch, total):
            self.output.reset(numpy.zeros((batch, total), dtype=numpy.float32))
        self._backend_init_()
        return None

    def numpy_run(self):
        self.output.map_invalidate()
        n = self.output.shape[0] if self.minibatch_size is None else int(self.minibatch_size)
        for inp, off, ln in zip(self.inputs, self.offsets, self.lengths):
            inp.map_read()
            self.output.mem[:n, off:off + ln] = inp.mem[:n].reshape(n, ln)

    def b200_init(self):
        self.init_vectors(self.output, *self.inputs)

    def b200_run(self):
        from .ops import kernels
        out = self.output.devmem
        for inp, off, ln in zip(self.inputs, self.offsets, self.lengths):
            inp.unmap()
            src = inp.devmem
            kernels.copy_2d(src.reshape(src.shape[0], -1) if src.dim() != 2 else src, out[:, off:off + ln])
        self.output.mark_device_written()
