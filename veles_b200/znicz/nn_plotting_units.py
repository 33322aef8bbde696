"""Neural-network plotters of the Znicz engine: ``Weights2D`` (a layer's kernels / neurons drawn as image tiles),
``SimilarWeights2D`` (the most correlated kernel pairs: dead or duplicated filters), ``MSEHistogram`` (distribution of
the per-sample errors of an MSE evaluator).  The reference's Znicz package is absent from the checkout; the names and
roles come from ``docs/source/manualrst_veles_workflow_creation.rst:550-644`` (``link_weights_plotter``,
``link_similar_weights_plotter``, ``link_mse_plotter``).  Headless: they render SVG files like every other plotter."""
from __future__ import annotations

import numpy

from ..plotter import Plotter
from ..plotting_units import _value
from ..svg import Canvas


def _tiles(weights, shape_hint=None, limit=64):
    """[n, features] -> list of 2-D (or HxWx3) arrays, one per neuron / kernel."""
    w = numpy.asarray(weights, dtype=numpy.float32)
    w = w.reshape(w.shape[0], -1)[:limit]
    out = []
    for row in w:
        if shape_hint is not None and int(numpy.prod(shape_hint)) == row.size:
            img = row.reshape(shape_hint)
            if img.ndim == 3 and img.shape[2] not in (1, 3):
                img = img.mean(axis=2)
            elif img.ndim == 3 and img.shape[2] == 1:
                img = img[:, :, 0]
        else:
            side = int(numpy.sqrt(row.size))
            img = row[:side * side].reshape(side, side) if side > 0 else row.reshape(1, -1)
        out.append(img)
    return out


class Weights2D(Plotter):
    """Tiles of the first ``limit`` rows of ``input.<input_field>`` (a weight matrix)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(Weights2D, self).__init__(workflow, **kwargs)
        self.input = None
        self.input_field = kwargs.get("input_field", "mem")
        self.limit = kwargs.get("limit", 64)
        self.get_shape_from = None       # optional [ky, kx, channels] of a convolutional layer
        self.transposed = kwargs.get("transposed", False)
        self.images = []

    def fill(self):
        w = numpy.asarray(_value(self.input, self.input_field if self.input_field != "mem" else None))
        if self.transposed:
            w = w.T
        self.images = _tiles(w, self.get_shape_from, self.limit)

    def redraw(self):
        n = max(len(self.images), 1)
        cols = int(numpy.ceil(numpy.sqrt(n)))
        c = Canvas(title=self.name, width=20 + 70 * cols, height=50 + 70 * ((n + cols - 1) // cols))
        for i, img in enumerate(self.images):
            c.image(img, 10 + 70 * (i % cols), 30 + 70 * (i // cols), 64, 64)
        return c.save(self.file_path())


class SimilarWeights2D(Weights2D):
    """Pairs of kernels whose normalised correlation exceeds ``form_threshold`` - drawn side by side."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(SimilarWeights2D, self).__init__(workflow, **kwargs)
        self.form_threshold = kwargs.get("form_threshold", 1.1)
        self.peak_threshold = kwargs.get("peak_threshold", 0.5)
        self.magnitude_threshold = kwargs.get("magnitude_threshold", 0.65)
        self.pairs = []

    def fill(self):
        w = numpy.asarray(_value(self.input, self.input_field if self.input_field != "mem" else None),
                          dtype=numpy.float32)
        if self.transposed:
            w = w.T
        w2 = w.reshape(w.shape[0], -1)
        c = w2 - w2.mean(axis=1, keepdims=True)
        norm = numpy.sqrt((c * c).sum(axis=1, keepdims=True))
        c = c / numpy.maximum(norm, 1e-12)
        corr = c.dot(c.T)
        numpy.fill_diagonal(corr, 0.0)
        thr = min(self.magnitude_threshold, 0.999)
        idx = numpy.argwhere(numpy.triu(corr, 1) > thr)
        order = numpy.argsort(-corr[idx[:, 0], idx[:, 1]]) if len(idx) else []
        self.pairs = [(int(idx[o, 0]), int(idx[o, 1]), float(corr[idx[o, 0], idx[o, 1]])) for o in order][:self.limit // 2]
        rows = []
        for a, b, _ in self.pairs:
            rows += [w2[a], w2[b]]
        self.images = _tiles(numpy.array(rows), self.get_shape_from, self.limit) if rows else []


class MSEHistogram(Plotter):
    """Histogram of the per-sample errors ``mse`` of an MSE evaluator."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(MSEHistogram, self).__init__(workflow, **kwargs)
        self.mse = None
        self.n_bars = kwargs.get("n_bars", 35)
        self.counts, self.edges = [], []

    def fill(self):
        v = numpy.asarray(_value(self, "mse"), dtype=numpy.float64).ravel()
        self.counts, self.edges = numpy.histogram(v, bins=self.n_bars)

    def redraw(self):
        c = Canvas(title=self.name)
        if len(self.counts):
            c.axes(self.edges[0], self.edges[-1], 0, max(max(self.counts), 1), xlabel="mse")
            c.bars(self.edges, self.counts)
        return c.save(self.file_path())
