"""Plotting units (reference: ``veles/plotting_units.py:52-903``): AccumulatingPlotter (curves over epochs),
MatrixPlotter (confusion matrix), ImagePlotter, ImmediatePlotter, Histogram, AutoHistogramPlotter,
MultiHistogram, TableMaxMin, SlaveStats.  All render headlessly to SVG through ``veles_b200.svg``."""
from __future__ import annotations

import numpy

from .memory import Array
from .plotter import Plotter
from .svg import Canvas


def _value(obj, field, offset=None):
    v = getattr(obj, field) if field is not None else obj
    if isinstance(v, Array):
        v.map_read()
        v = v.mem
    if offset is not None:
        v = v[offset]
    return v


class AccumulatingPlotter(Plotter):
    """Appends ``input.<input_field>[input_offset]`` every run and draws the curve."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(AccumulatingPlotter, self).__init__(workflow, **kwargs)
        self.plot_style = kwargs.get("plot_style", "k-")
        self.label = kwargs.get("label", "")
        self.fit_poly_power = kwargs.get("fit_poly_power", 0)
        self.minimum, self.maximum = kwargs.get("minimum"), kwargs.get("maximum")
        self.input = None
        self.input_field = kwargs.get("input_field")
        self.input_offset = kwargs.get("input_offset")
        self.values = []
        self.redraw_threshold = 0

    def fill(self):
        v = _value(self.input, self.input_field, self.input_offset)
        self.values.append(float(v))

    def redraw(self):
        c = Canvas(title=self.name)
        ys = self.values or [0.0]
        xs = list(range(len(ys)))
        lo = min(ys) if self.minimum is None else self.minimum
        hi = max(ys) if self.maximum is None else self.maximum
        c.axes(0, max(len(ys) - 1, 1), lo, hi, xlabel="run", ylabel=self.label)
        c.line(xs, ys, label=self.label or None)
        if self.fit_poly_power and len(ys) > self.fit_poly_power:
            coef = numpy.polyfit(xs, ys, self.fit_poly_power)
            c.line(xs, numpy.polyval(coef, xs), index=1, label="fit")
        return c.save(self.file_path())


class MatrixPlotter(Plotter):
    """Confusion matrix (``input.<input_field>`` is a square count matrix)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(MatrixPlotter, self).__init__(workflow, **kwargs)
        self.input = None
        self.input_field = kwargs.get("input_field")
        self.reversed_labels_mapping = kwargs.get("reversed_labels_mapping")
        self.matrix = None

    def fill(self):
        m = _value(self.input, self.input_field)
        self.matrix = None if m is None else numpy.array(m)

    def redraw(self):
        c = Canvas(title=self.name, width=520, height=520)
        if self.matrix is not None:
            labels = [str(l) for l in (self.reversed_labels_mapping or range(self.matrix.shape[1]))]
            c.matrix(self.matrix, labels)
        return c.save(self.file_path())


class ImagePlotter(Plotter):
    """Draws up to ``count`` samples from the linked arrays as images."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(ImagePlotter, self).__init__(workflow, **kwargs)
        self.inputs = []
        self.input_fields = []
        self.count = kwargs.get("count", 4)
        self.images = []

    def fill(self):
        self.images = []
        for obj, field in zip(self.inputs, self.input_fields):
            v = _value(obj, field)
            for i in range(min(self.count, len(v))):
                img = numpy.array(v[i])
                if img.ndim == 1:
                    side = int(round(img.size ** 0.5))
                    img = img[:side * side].reshape(side, side)
                self.images.append(img)

    def redraw(self):
        n = max(len(self.images), 1)
        c = Canvas(title=self.name, width=20 + 130 * min(n, 4), height=50 + 130 * ((n + 3) // 4))
        for i, img in enumerate(self.images):
            c.image(img, 10 + 130 * (i % 4), 30 + 130 * (i // 4), 120, 120)
        return c.save(self.file_path())


class ImmediatePlotter(Plotter):
    """Plots whole vectors (one curve per input) as they are at the moment of the run."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(ImmediatePlotter, self).__init__(workflow, **kwargs)
        self.inputs, self.input_fields, self.input_styles = [], [], []
        self.curves = []

    def fill(self):
        self.curves = [numpy.array(_value(o, f), dtype=numpy.float64).ravel()
                       for o, f in zip(self.inputs, self.input_fields)]

    def redraw(self):
        c = Canvas(title=self.name)
        if self.curves:
            lo = min(float(v.min()) for v in self.curves)
            hi = max(float(v.max()) for v in self.curves)
            c.axes(0, max(max(len(v) for v in self.curves) - 1, 1), lo, hi)
            for i, v in enumerate(self.curves):
                c.line(range(len(v)), v, index=i)
        return c.save(self.file_path())


class Histogram(Plotter):
    """Bar chart of precomputed ``x`` (bin edges) / ``y`` (counts)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(Histogram, self).__init__(workflow, **kwargs)
        self.x, self.y = None, None
        self.x_copy, self.y_copy = [], []

    def fill(self):
        self.x_copy = list(numpy.ravel(_value(self, "x")))
        self.y_copy = list(numpy.ravel(_value(self, "y")))

    def redraw(self):
        c = Canvas(title=self.name)
        if self.y_copy:
            edges = self.x_copy if len(self.x_copy) == len(self.y_copy) + 1 else list(range(len(self.y_copy) + 1))
            c.axes(edges[0], edges[-1], 0, max(self.y_copy))
            c.bars(edges, self.y_copy)
        return c.save(self.file_path())


class AutoHistogramPlotter(Plotter):
    """Histogram of all values of a linked array (e.g. a layer's weights)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(AutoHistogramPlotter, self).__init__(workflow, **kwargs)
        self.input = None
        self.input_field = kwargs.get("input_field")
        self.n_bars = kwargs.get("n_bars", 40)
        self.counts, self.edges = [], []

    def fill(self):
        v = numpy.asarray(_value(self.input, self.input_field), dtype=numpy.float64).ravel()
        self.counts, self.edges = numpy.histogram(v, bins=self.n_bars)

    def redraw(self):
        c = Canvas(title=self.name)
        if len(self.counts):
            c.axes(self.edges[0], self.edges[-1], 0, max(self.counts))
            c.bars(self.edges, self.counts)
        return c.save(self.file_path())


class MultiHistogram(Plotter):
    """One small histogram per row of the input (e.g. per neuron weights)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(MultiHistogram, self).__init__(workflow, **kwargs)
        self.input = None
        self.input_field = kwargs.get("input_field")
        self.hist_number = kwargs.get("hist_number", 8)
        self.n_bars = kwargs.get("n_bars", 20)
        self.hists = []

    def fill(self):
        v = numpy.asarray(_value(self.input, self.input_field), dtype=numpy.float64)
        v = v.reshape(v.shape[0], -1)
        self.hists = [numpy.histogram(v[i], bins=self.n_bars) for i in range(min(self.hist_number, len(v)))]

    def redraw(self):
        c = Canvas(title=self.name, height=60 + 90 * max(len(self.hists), 1))
        for i, (counts, edges) in enumerate(self.hists):
            c.plot = (56, 30 + 90 * i, c.w - 20, 100 + 90 * i)
            c.axes(edges[0], edges[-1], 0, max(counts.max(), 1))
            c.bars(edges, counts)
        return c.save(self.file_path())


class TableMaxMin(Plotter):
    """Table of max / min of several arrays."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(TableMaxMin, self).__init__(workflow, **kwargs)
        self.y, self.col_labels = [], ["max", "min"]
        self.row_labels = []
        self.values = []

    def fill(self):
        self.values = []
        for arr in self.y:
            v = numpy.asarray(_value(arr, None), dtype=numpy.float64)
            self.values.append((float(v.max()), float(v.min())))

    def redraw(self):
        c = Canvas(title=self.name, height=60 + 18 * max(len(self.values), 1))
        for i, (mx, mn) in enumerate(self.values):
            label = self.row_labels[i] if i < len(self.row_labels) else "array %d" % i
            c.text(20, 50 + 18 * i, "%-24s max %.6g   min %.6g" % (label, mx, mn))
        return c.save(self.file_path())


class SlaveStats(Plotter):
    """Jobs done per slave (master side)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(SlaveStats, self).__init__(workflow, **kwargs)
        self.stats = {}

    def fill(self):
        agent = getattr(self.launcher, "agent", None)
        self.stats = agent.slaves_status() if agent is not None and hasattr(agent, "slaves_status") else {}

    def redraw(self):
        c = Canvas(title=self.name, height=60 + 18 * max(len(self.stats), 1))
        for i, (sid, st) in enumerate(sorted(self.stats.items())):
            c.text(20, 50 + 18 * i, "%s  %s  jobs=%s  power=%s" % (sid, st.get("state"), st.get("jobs"), st.get("power")))
        return c.save(self.file_path())
