"""Publisher unit: at the end of training gathers results, configuration, seeds, unit timings, plots and the
workflow graph and renders them with the selected backends (reference: ``veles/publishing/publisher.py``)."""
from __future__ import annotations

import time

from .. import prng
from ..config import root, Config, get
from ..distributable import TriviallyDistributable
from ..units import Unit
from .backend import BackendRegistry


class Publisher(Unit, TriviallyDistributable):
    """Collects the facts of a finished run (metrics, timings, configuration, plots, graph) and renders them
    through the configured report back-ends."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "SERVICE")
        super(Publisher, self).__init__(workflow, **kwargs)
        self.backends = dict(kwargs.get("backends", {"markdown": {}}))
        self.templates = kwargs.get("templates")
        self.reports = []
        self.loader_unit = kwargs.get("loader_unit")

    def initialize(self, **kwargs):
        for name in self.backends:
            if name not in BackendRegistry.backends:
                raise ValueError("Unknown publishing backend %r; known: %s" % (name, sorted(BackendRegistry.backends)))

    def gather_info(self):
        wf = self.workflow
        launcher = self.launcher
        cfg = {name: node.to_dict() for name, node in root.__content__.items()
               if name != "common" and isinstance(node, Config)}
        plots = [u.last_file for u in wf.units if getattr(u, "last_file", None)]
        graphics = getattr(launcher, "graphics", None)
        if graphics is not None:
            graphics.flush()
            plots += list(getattr(graphics, "rendered", []))
        return {"name": wf.name, "description": (type(wf).__doc__ or "").strip(), "id": getattr(launcher, "id", None),
                "mode": getattr(launcher, "mode", "standalone"),
                "device": repr(getattr(getattr(wf, "device", None), "device_info", None)),
                "run_time": time.time() - (getattr(launcher, "start_time", None) or time.time()),
                "epochs": getattr(getattr(wf, "loader", None), "epoch_number", None),
                "seeds": {k: repr(g.seed_value)[:80] for k, g in prng.random_generator._generators.items()},
                "results": wf.gather_results(), "config": cfg,
                "unit_times": sorted(((u.name, u.total_run_time) for u in wf.units), key=lambda kv: -kv[1])[:10],
                "plots": sorted(set(p for p in plots if p)), "graph": wf.generate_graph(write_on_disk=False)}

    def run(self):
        if get(root.common.disable.publishing, False):
            return
        info = self.gather_info()
        self.reports = []
        for name, kwargs in self.backends.items():
            path = BackendRegistry.backends[name](**(kwargs or {})).render(info)
            self.reports.append(path)
            self.info("Published the %s report to %s", name, path)
