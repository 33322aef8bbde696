"""ResizableAll2All (``all2all_resizable``): a fully connected layer whose number of neurons can be changed
between runs; existing weights are kept, new rows are initialised like fresh ones."""
from __future__ import annotations

import numpy

from .all2all import All2All
from .gd import GradientDescent


class ResizableAll2All(All2All):
    """Fully connected layer whose number of outputs can be changed between runs (weights are extended or
    cropped)."""
    TRANSPOSED_OK = False
    MAPPING = {"all2all_resizable"}

    def resize(self, n_out):
        """Change the number of output neurons; returns True if anything changed."""
        old = self.neurons_number
        if n_out == old:
            return False
        self.weights.map_read()
        self.bias.map_read()
        w_old, b_old = self.weights.mem, self.bias.mem if self.bias else None
        n_in = w_old.shape[1]
        w = numpy.zeros((n_out, n_in), dtype=numpy.float32)
        self.fill_array(self.weights_filling, w, self.weights_stddev)
        keep = min(old, n_out)
        w[:keep] = w_old[:keep]
        self.output_sample_shape = (n_out,)
        self.weights.reset(w)
        if b_old is not None:
            b = numpy.zeros(n_out, dtype=numpy.float32)
            self.fill_array(self.bias_filling, b, self.bias_stddev)
            b[:keep] = b_old[:keep]
            self.bias.reset(b)
        self.output.reset()
        self.weights_bf16_ = None
        self._is_initialized = False
        return True


class GDResizableAll2All(GradientDescent):
    """GradientDescent for ResizableAll2All."""
    MAPPING = {"all2all_resizable"}
