"""Device random-bytes unit: xorshift1024* streams, one per state, interleaved.

Reference: ``veles/prng/uniform.py:49-175`` (AcceleratedUnit around ``cuda/random.cu``): ``num_states``
(default 256) x 16 u64 of state; every round each state emits 16 u64, laid out
``output[round*16*S + i*S + state]``.  The device path is ``vb_xorshift1024star`` (csrc/prng.cu), one
thread per state with the 16-word state in registers; the numpy oracle is
``prng.random_generator.xorshift1024star_fill`` and both produce the same bit stream.
"""
from __future__ import annotations

import numpy

from ..accelerated_units import AcceleratedUnit
from ..memory import Array
from . import random_generator


class Uniform(AcceleratedUnit):
    """Fills ``output`` (raw bytes, uint64 granularity) with xorshift1024* draws every run."""

    def __init__(self, workflow, **kwargs):
        super(Uniform, self).__init__(workflow, **kwargs)
        self.num_states = int(kwargs.get("num_states", 256))
        self.output_bytes = int(kwargs.get("output_bytes", 0))
        self.prng = kwargs.get("prng", random_generator.get())
        self.states = Array()
        self.output = Array()

    @property
    def rounds(self):
        per_round = self.num_states * 16 * 8
        return max(1, (self.output_bytes + per_round - 1) // per_round)

    def initialize(self, device=None, **kwargs):
        self.device = device
        if not self.states or self.states.shape[0] != self.num_states:
            seed = self.prng.randint(1, 1 << 62, size=self.num_states * 16)
            self.states.reset(numpy.asarray(seed, dtype=numpy.uint64).reshape(self.num_states, 16))
        n = self.rounds * 16 * self.num_states
        if not self.output or self.output.size != n:
            self.output.reset(numpy.zeros(n, dtype=numpy.uint64))
        self._backend_init_()
        return None

    # -- numpy oracle -------------------------------------------------------------------------------------
    def numpy_init(self):
        pass

    def numpy_run(self):
        self.states.map_write()
        self.output.map_invalidate()
        self.output.mem[:] = random_generator.xorshift1024star_fill(self.states.mem, self.rounds)

    def numpy_fill(self, nbytes):
        """Reference-compatible helper: advance the host states and return ``nbytes`` of output."""
        rounds = max(1, (nbytes + self.num_states * 128 - 1) // (self.num_states * 128))
        self.states.map_write()
        return random_generator.xorshift1024star_fill(self.states.mem, rounds).view(numpy.uint8)[:nbytes]

    # -- device -------------------------------------------------------------------------------------------
    def b200_init(self):
        import torch
        dev = self.device_
        self.states_dev_ = torch.from_numpy(self.states.mem.view(numpy.int64).copy()).to(dev.torch_device)
        self.output.initialize(dev, storage=torch.zeros(self.output.size, dtype=torch.int64, device=dev.torch_device))

    def b200_run(self):
        from ..ops import kernels
        kernels.xorshift1024star(self.states_dev_, self.rounds, self.output.devmem)
        self.output.mark_device_written()

    def sync_states(self):
        """Device states -> host Array (before pickling / comparing with the oracle)."""
        if getattr(self, "states_dev_", None) is not None:
            self.states.mem[...] = self.states_dev_.cpu().numpy().view(numpy.uint64).reshape(self.states.shape)
