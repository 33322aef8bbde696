"""Plotter base unit: a picklable data holder whose ``redraw()`` renders headlessly.

Parity target: ``veles/plotter.py:47-179``: ``run()`` -> ``fill()`` (copy what must be drawn) -> stripped pickle
-> graphics server; the client calls ``redraw()``.  The reference renders with matplotlib GUI backends in a
separate process; no graphics stack exists on the B200 boxes, so ``redraw()`` writes an SVG file (pure text,
no dependency) — or uses matplotlib's Agg backend when matplotlib is importable."""
from __future__ import annotations

import os
import time

from .config import root, get
from .distributable import TriviallyDistributable
from .units import Unit


class Plotter(Unit, TriviallyDistributable):
    hide_from_registry = True
    server_shutdown_registered = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "PLOTTER")
        super(Plotter, self).__init__(workflow, **kwargs)
        self.redraw_threshold = kwargs.get("redraw_threshold", 0.5)
        self.clear_plot = kwargs.get("clear_plot", False)
        self.redraw_plot = kwargs.get("redraw_plot", True)
        self.output_dir = kwargs.get("output_dir", get(root.common.graphics.dir, "plots"))
        self.last_file = None
        self._last_run_ = 0.0

    def init_unpickled(self):
        super(Plotter, self).init_unpickled()
        self._last_run_ = 0.0

    @property
    def graphics_server(self):
        launcher = self.launcher
        return getattr(launcher, "graphics", None) if launcher is not None else None

    @property
    def enabled(self):
        wf = self.workflow
        return bool(getattr(wf, "plotters_are_enabled", True)) and not get(root.common.disable.plotting, False)

    def initialize(self, **kwargs):
        pass

    def fill(self):
        """Copy the data to plot from the linked attributes (override)."""

    def redraw(self):
        """Render; returns the path of the written file (override)."""
        return None

    def run(self):
        if not self.enabled:
            return
        now = time.time()
        if now - self._last_run_ < self.redraw_threshold and self._last_run_:
            return
        self._last_run_ = now
        self.fill()
        server = self.graphics_server
        if server is not None:
            server.enqueue(self)
        else:
            self.last_file = self.redraw()

    def file_path(self, ext="svg"):
        os.makedirs(self.output_dir, exist_ok=True)
        safe = "".join(c if c.isalnum() or c in "-_" else "_" for c in self.name)
        return os.path.join(self.output_dir, "%s.%s" % (safe, ext))
