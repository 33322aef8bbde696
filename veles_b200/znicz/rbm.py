"""Restricted Boltzmann Machine trained with one step of contrastive divergence (CD-1) — named among the Znicz
algorithms (``docs/source/manualrst_veles_algorithms.rst:113-135``, "created, not tested" there).  torch/host
implementation kept off the benchmarked path."""
from __future__ import annotations

import numpy

from ..accelerated_units import AcceleratedUnit
from ..distributable import TriviallyDistributable
from ..memory import Array
from .. import prng


class RBM(AcceleratedUnit, TriviallyDistributable):
    """Restricted Boltzmann machine layer trained by contrastive divergence."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(RBM, self).__init__(workflow, **kwargs)
        self.n_hidden = int(kwargs.get("n_hidden", 64))
        self.learning_rate = kwargs.get("learning_rate", 0.05)
        self.rand = kwargs.get("rand", prng.get())
        self.weights, self.vbias, self.hbias = Array(), Array(), Array()
        self.output = Array(shallow_pickle=True)
        self.reconstruction_error = 0.0
        self.demand("input")

    def initialize(self, device=None, **kwargs):
        if self.input is None or not self.input:
            return True
        self.device = device
        nv = self.input.sample_size
        if not self.weights:
            w = numpy.zeros((self.n_hidden, nv), dtype=numpy.float32)
            self.rand.fill_normal_real(w, 0, 0.01)
            self.weights.reset(w)
            self.vbias.reset(numpy.zeros(nv, dtype=numpy.float32))
            self.hbias.reset(numpy.zeros(self.n_hidden, dtype=numpy.float32))
        self.output.reset(numpy.zeros((self.input.shape[0], self.n_hidden), dtype=numpy.float32))
        return None

    @staticmethod
    def _sig(x):
        return 1.0 / (1.0 + numpy.exp(-x))

    def numpy_init(self):
        pass

    b200_init = numpy_init

    def numpy_run(self):
        self.input.map_read()
        v0 = self.input.mem.reshape(self.input.shape[0], -1).astype(numpy.float32)
        w, vb, hb = self.weights.mem, self.vbias.mem, self.hbias.mem
        h0 = self._sig(v0.dot(w.T) + hb)
        hs = (self.rand.random_sample(h0.shape) < h0).astype(numpy.float32)
        v1 = self._sig(hs.dot(w) + vb)
        h1 = self._sig(v1.dot(w.T) + hb)
        n = v0.shape[0]
        w += self.learning_rate * (h0.T.dot(v0) - h1.T.dot(v1)) / n
        vb += self.learning_rate * (v0 - v1).mean(axis=0)
        hb += self.learning_rate * (h0 - h1).mean(axis=0)
        self.reconstruction_error = float(((v0 - v1) ** 2).mean())
        self.output.map_invalidate()
        self.output.mem[...] = h0

    b200_run = numpy_run
