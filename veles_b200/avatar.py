"""Avatar: clones selected attributes/Arrays of another unit each run, so that unit (typically a Loader)
can already prepare the NEXT minibatch while the network trains on the copy
(reference: ``veles/avatar.py:22-129``, device-to-device copies at ``:112-123``).
On the B200 path the copies are ``cudaMemcpyAsync`` D2D on the current stream; ``StreamLoader``
double-buffers host uploads the same way."""
from __future__ import annotations

import numpy

from .accelerated_units import AcceleratedUnit
from .distributable import TriviallyDistributable
from .memory import Array
from .mutable import Bool


class Avatar(AcceleratedUnit, TriviallyDistributable):
    """Mirror of another unit's attributes: clones the listed Arrays / values on every run, so that a reader (a
    plotter, a saver) sees a consistent copy while the original keeps changing (reference: veles/avatar.py)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "LOADER")
        super(Avatar, self).__init__(workflow, **kwargs)
        self.reals = {}     # unit -> [attribute names]
        self.vectors = {}   # real Array id -> (real, clone)

    def clone(self):
        """Create the mirror attributes on self (call after ``reals`` is filled)."""
        for unit, attrs in self.reals.items():
            for attr in attrs:
                value = getattr(unit, attr)
                if isinstance(value, Array):
                    clone = Array(shallow_pickle=True)
                    if value:
                        clone.reset(value.mem.copy())
                    setattr(self, attr, clone)
                    self.vectors[id(value)] = (value, clone)
                elif isinstance(value, Bool):
                    setattr(self, attr, Bool(bool(value)))
                else:
                    setattr(self, attr, value)

    def initialize(self, device=None, **kwargs):
        self.device = device
        for real, clone in self.vectors.values():
            if not real:
                return True
            if not clone or clone.shape != real.shape:
                clone.reset(numpy.zeros(real.shape, dtype=real.dtype))
            if self.on_device and real.devmem is not None:
                clone.dev_dtype = real.devmem.dtype
                clone.initialize(self.device_)
        return None

    def numpy_run(self):
        for unit, attrs in self.reals.items():
            for attr in attrs:
                value = getattr(unit, attr)
                if isinstance(value, Array):
                    real, clone = self.vectors[id(value)]
                    if real.devmem is not None and clone.devmem is not None:
                        clone.devmem.copy_(real.devmem, non_blocking=True)
                        clone.mark_device_written()
                    else:
                        real.map_read()
                        clone.map_invalidate()
                        numpy.copyto(clone.mem, real.mem)
                elif isinstance(value, Bool):
                    getattr(self, attr).__ilshift__(bool(value))
                else:
                    setattr(self, attr, value)

    b200_run = numpy_run
