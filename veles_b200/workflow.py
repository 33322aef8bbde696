"""Workflow — container of units, itself a unit.

Parity target: ``veles/workflow.py:87-1051``: container protocol, BFS ``initialize`` with a
retry queue (:303-349), ``run`` = ``start_point.run_dependent()`` + wait for the finish event
(:351-369), ``on_workflow_finished`` (:377-400), distributed fan-out of the four data methods
in dependency order (:456-548), ``drop_slave``, ``do_job`` (:558-573), ``generate_graph`` DOT
export (:628-765), ``print_stats`` (:767-825), ``gather_results``/``write_results`` (:827-849),
``checksum`` (:851-866), ``package_export`` (:868-975), ``change_unit`` (:977-1051).
"""
from __future__ import annotations

import hashlib
import inspect
import io
import json
import os
import sys
import tarfile
import threading
import time
import zipfile
from collections import OrderedDict

import numpy

from .config import root, get
from .error import VelesException
from .plumbing import StartPoint, EndPoint, Repeater
from .result_provider import IResultProvider
from .thread_pool import ThreadPool
from .units import Unit, Container, TrivialUnit
from .verified import interface_methods


class NoMoreJobs(Exception):
    """Raised by a unit's generate_data_for_slave when the master has nothing left."""


class MultiMap(OrderedDict):
    """name -> [units] with insertion order."""

    def add(self, key, value):
        self.setdefault(key, []).append(value)

    def remove(self, key, value):
        lst = self.get(key)
        if lst and value in lst:
            lst.remove(value)
            if not lst:
                del self[key]


class Workflow(Container):
    """Base class of unit graphs."""
    hide_from_registry = True
    VIEW_GROUP_COLORS = {"PLUMBING": "gray", "WORKER": "lightblue", "LOADER": "lightyellow",
                         "TRAINER": "coral", "EVALUATOR": "plum", "START_END": "lightgrey",
                         "PLOTTER": "palegreen", "SERVICE": "khaki"}

    def __init__(self, workflow, **kwargs):
        self._plotters_are_enabled = kwargs.get("enable_plotters",
                                                not get(root.common.disable.plotting, False))
        self._sync = kwargs.get("sync", True)
        self._units = []
        self._result_file = kwargs.get("result_file")
        self._restored_from_snapshot_ = None
        self._finished_ = False
        self.debug_deadlocks = False
        super(Workflow, self).__init__(workflow, **kwargs)
        self.start_point = StartPoint(self)
        self.end_point = EndPoint(self)
        self.negotiates_on_connect = True
        self._is_running = False
        self._run_time_ = 0.0

    def init_unpickled(self):
        super(Workflow, self).init_unpickled()
        self._sync_event_ = threading.Event()
        self._sync_event_.set()
        self._run_time_started_ = time.time()
        self._run_time_ = 0.0
        self._finished_ = False
        self._do_job_callback_ = None
        self._stop_lock_ = threading.Lock()
        self.__dict__.setdefault("_restored_from_snapshot_", None)
        self._own_pool_ = None
        self._method_time_ = {}
        self.units_by_name_ = None

    def __getstate__(self):
        state = super(Workflow, self).__getstate__()
        # the launcher (or any non-picklable host) is re-attached after loading
        host = state.get("_workflow")
        if host is not None and not isinstance(host, Unit):
            state["_workflow"] = None
        return state

    def __setstate__(self, state):
        super(Workflow, self).__setstate__(state)
        self._restored_from_snapshot_ = True

    # -- hosting --------------------------------------------------------------------------------
    @property
    def workflow(self):
        return self._workflow

    @workflow.setter
    def workflow(self, value):
        # a Workflow may be hosted by a launcher-like object or another workflow
        if value is None:
            raise ValueError("Workflow must have a launcher or a parent workflow")
        if self._workflow is not None and hasattr(self._workflow, "del_ref"):
            self._workflow.del_ref(self)
        self._workflow = value
        if hasattr(value, "add_ref"):
            value.add_ref(self)

    @property
    def launcher(self):
        host = self._workflow
        while isinstance(host, Workflow):
            host = host._workflow
        return host

    @property
    def thread_pool(self):
        host = self._workflow
        pool = getattr(host, "thread_pool", None) if host is not None else None
        if pool is None:
            if self._own_pool_ is None:
                tp = root.common.engine.thread_pool
                self._own_pool_ = ThreadPool(minthreads=get(tp.minthreads, 2), maxthreads=get(tp.maxthreads, 2),
                                             name=self.name)
            pool = self._own_pool_
        return pool

    def _host_attr(self, name, default):
        host = self._workflow
        return getattr(host, name, default) if host is not None else default

    @property
    def is_master(self):
        return bool(self._host_attr("is_master", False))

    @property
    def is_slave(self):
        return bool(self._host_attr("is_slave", False))

    @property
    def is_standalone(self):
        return bool(self._host_attr("is_standalone", True))

    @property
    def is_main(self):
        return not isinstance(self._workflow, Workflow)

    @property
    def interactive(self):
        return bool(self._host_attr("interactive", False))

    @property
    def testing(self):
        return bool(self._host_attr("testing", False))

    @property
    def restored_from_snapshot(self):
        if self._restored_from_snapshot_ is not None:
            return self._restored_from_snapshot_
        host = self._workflow
        return bool(getattr(host, "restored_from_snapshot", False)) if isinstance(host, Workflow) else False

    @property
    def is_running(self):
        return self._is_running

    @property
    def plotters_are_enabled(self):
        return self._plotters_are_enabled

    @plotters_are_enabled.setter
    def plotters_are_enabled(self, value):
        self._plotters_are_enabled = bool(value)

    @property
    def result_file(self):
        return self._result_file

    @result_file.setter
    def result_file(self, value):
        self._result_file = value

    # -- container protocol -----------------------------------------------------------------------
    def add_ref(self, unit):
        if unit is self:
            raise ValueError("Attempted to add self to self")
        if unit not in self._units:
            self._units.append(unit)
        self.units_by_name_ = None

    def del_ref(self, unit):
        if unit in self._units:
            self._units.remove(unit)
        self.units_by_name_ = None

    @property
    def units(self):
        return list(self._units)

    @property
    def units_in_dependency_order(self):
        order = [u for u in self.start_point.dependent_units()]
        seen = set(id(u) for u in order)
        return order + [u for u in self._units if id(u) not in seen]

    def index_of(self, unit):
        return self._units.index(unit)

    def _by_name(self):
        if self.units_by_name_ is None:
            mm = MultiMap()
            for u in self._units:
                mm.add(u.name, u)
            self.units_by_name_ = mm
        return self.units_by_name_

    def __getitem__(self, key):
        if isinstance(key, str):
            found = self._by_name().get(key)
            if not found:
                raise KeyError(key)
            return found[0] if len(found) == 1 else list(found)
        return self._units[key]

    def __iter__(self):
        return iter(list(self._units))

    def __len__(self):
        return len(self._units)

    def __contains__(self, unit):
        return unit in self._units

    # -- life cycle -----------------------------------------------------------------------------------
    def initialize(self, **kwargs):
        """Initialize units in BFS order from start_point; units returning True are retried."""
        self._finished_ = False
        self.event("initialize", "begin")
        snapshot = kwargs.get("snapshot", False)
        units = self.units_in_dependency_order
        fin_text = "all units are initialized"
        pending = list(units)
        max_rounds = len(pending) + 2
        rounds = 0
        while pending:
            postponed = []
            for unit in pending:
                if unit is self:
                    continue
                try:
                    retry = unit.initialize(**kwargs)
                except Exception:
                    self.error("Unit \"%s\" failed to initialize", unit.name)
                    raise
                if retry:
                    postponed.append(unit)
            if len(postponed) == len(pending):
                rounds += 1
                if rounds > max_rounds:
                    raise VelesException("Units %s keep postponing their initialization"
                                         % ", ".join(u.name for u in postponed))
            pending = postponed
        if self.restored_from_snapshot or snapshot:
            # gates may have been pickled half-open: start from a clean state
            for unit in units:
                if unit is not self:
                    unit.close_gate()
        self.debug(fin_text)
        self.event("initialize", "end")
        self._is_initialized = True
        return None

    def run(self):
        """Start the graph and (by default) block until EndPoint fires."""
        self._sync_event_.clear()
        self._finished_ = False
        self._is_running = True
        self._stopped = False
        for unit in self._units:
            unit.stopped = False  # a finished workflow stopped its units; a new run (next slave job) revives them
        self._run_time_started_ = time.time()
        self.event("run", "begin")
        if not self.is_master:
            # a nested workflow runs inside its parent's dispatch loop: give it a loop of its own,
            # otherwise its first unit would only be queued behind this very call
            from . import units as _units
            saved = getattr(_units._tls, "queue", None)
            _units._tls.queue = None
            try:
                self.start_point.run_dependent()
            finally:
                _units._tls.queue = saved
        if self._sync:
            self.wait_finished()

    def wait_finished(self, timeout=None):
        step = 0.2
        waited = 0.0
        while not self._sync_event_.wait(step):
            waited += step
            pool = self.thread_pool
            if pool is not None and pool.failure is not None:
                self._is_running = False
                self._sync_event_.set()
                raise pool.failure
            if timeout is not None and waited >= timeout:
                return False
        pool = self.thread_pool
        if pool is not None and pool.failure is not None:
            raise pool.failure
        return True

    def stop(self):
        """Ask every unit to stop and finish the workflow."""
        with self._stop_lock_:
            for unit in self._units:
                if unit is not self and not unit.stopped:
                    unit.stop()
        if self._is_running:
            self.on_workflow_finished()

    def on_workflow_finished(self):
        with self._stop_lock_:
            if self._finished_:
                return
            self._finished_ = True
        for unit in self._units:
            if not unit.stopped and unit is not self:
                try:
                    unit.stop()
                except Exception:  # noqa: BLE001 - a failing stop() must not mask the finish
                    self.exception("%s.stop() failed", unit.name)
        self._run_time_ += time.time() - self._run_time_started_
        self._is_running = False
        self.event("run", "end")
        if self.is_standalone or self.is_master:
            if self._result_file:
                self.write_results()
        host = self._workflow
        try:
            if self.is_slave and self._do_job_callback_ is not None:
                cb, self._do_job_callback_ = self._do_job_callback_, None
                cb(self.generate_data_for_master())
            elif host is not None and hasattr(host, "on_workflow_finished") and not isinstance(host, Workflow):
                host.on_workflow_finished()   # launchers only: a parent workflow just sees run() return
        finally:
            self._sync_event_.set()

    # -- distributed fan-out (job-level control plane) --------------------------------------------------
    def _timed(self, name, fn, *args):
        t0 = time.perf_counter()
        try:
            return fn(*args)
        finally:
            self._method_time_[name] = self._method_time_.get(name, 0.0) + time.perf_counter() - t0

    @staticmethod
    def _distributable(unit):
        return all(callable(getattr(unit, m, None)) for m in
                   ("generate_data_for_master", "generate_data_for_slave", "apply_data_from_master",
                    "apply_data_from_slave", "drop_slave"))

    def _dist_units(self):
        # creation order: identical on master and slave (both execute the same constructor), whereas the
        # BFS order depends on mode-specific links (e.g. where EndPoint hangs on a slave)
        return [u for u in self._units if u is not self and self._distributable(u)]

    def generate_data_for_master(self):
        self.event("generate_data", "begin", side="slave")
        data = self._timed("generate_data_for_master", lambda: [
            u.generate_data_for_master_threadsafe() if hasattr(u, "generate_data_for_master_threadsafe")
            else u.generate_data_for_master() for u in self._dist_units()])
        self.event("generate_data", "end", side="slave")
        return data

    def generate_data_for_slave(self, slave=None):
        """Master: job for ``slave`` or False when some unit has no data yet (sync point)."""
        units = self._dist_units()
        if not all(getattr(u, "has_data_for_slave", True) for u in units):
            return False
        self.event("generate_data", "begin", side="master", slave=str(getattr(slave, "id", slave)))
        data = []

        def gen():
            for u in units:
                data.append(u.generate_data_for_slave_threadsafe(slave)
                            if hasattr(u, "generate_data_for_slave_threadsafe")
                            else u.generate_data_for_slave(slave))
        try:
            self._timed("generate_data_for_slave", gen)
        except NoMoreJobs:
            self.on_workflow_finished()
            return None
        finally:
            self.event("generate_data", "end", side="master")
        return data

    def apply_data_from_master(self, data):
        units = self._dist_units()
        if not isinstance(data, (list, tuple)) or len(data) != len(units):
            raise ValueError("Job size %s does not match the number of distributable units %d"
                             % (len(data) if hasattr(data, "__len__") else "?", len(units)))
        self.event("apply_data", "begin", side="slave")

        def app():
            for u, d in zip(units, data):
                if d is not None:
                    (u.apply_data_from_master_threadsafe if hasattr(u, "apply_data_from_master_threadsafe")
                     else u.apply_data_from_master)(d)
        self._timed("apply_data_from_master", app)
        self.event("apply_data", "end", side="slave")

    def apply_data_from_slave(self, data, slave=None):
        units = self._dist_units()
        if not isinstance(data, (list, tuple)) or len(data) != len(units):
            raise ValueError("Update size does not match the number of distributable units")
        self.event("apply_data", "begin", side="master", slave=str(getattr(slave, "id", slave)))

        def app():
            for u, d in zip(units, data):
                if d is not None:
                    (u.apply_data_from_slave_threadsafe if hasattr(u, "apply_data_from_slave_threadsafe")
                     else u.apply_data_from_slave)(d, slave)
        self._timed("apply_data_from_slave", app)
        self.event("apply_data", "end", side="master")
        return True

    def drop_slave(self, slave=None):
        self.event("drop_slave", "single", slave=str(getattr(slave, "id", slave)))
        for u in self._dist_units():
            (u.drop_slave_threadsafe if hasattr(u, "drop_slave_threadsafe") else u.drop_slave)(slave)
        self.warning("Dropped the job from %s", getattr(slave, "id", slave))

    def do_job(self, data, update, callback):
        """Slave: apply a job, run one pass, hand the update to ``callback``."""
        self.apply_data_from_master(data)
        self._do_job_callback_ = callback
        try:
            self._sync = False
            self.run()
        except Exception:
            self._do_job_callback_ = None
            raise

    def generate_initial_data_for_master(self):
        return [getattr(u, "generate_initial_data_for_master", lambda: None)()
                if getattr(u, "negotiates_on_connect", False) else None for u in self._dist_units()]

    def generate_initial_data_for_slave(self, slave=None):
        return [getattr(u, "generate_initial_data_for_slave", lambda s: None)(slave)
                if getattr(u, "negotiates_on_connect", False) else None for u in self._dist_units()]

    def apply_initial_data_from_master(self, data):
        if not data:
            return
        for u, d in zip(self._dist_units(), data):
            if d is not None and hasattr(u, "apply_initial_data_from_master"):
                u.apply_initial_data_from_master(d)

    def apply_initial_data_from_slave(self, data, slave=None):
        if not data:
            return
        for u, d in zip(self._dist_units(), data):
            if d is not None and hasattr(u, "apply_initial_data_from_slave"):
                u.apply_initial_data_from_slave(d, slave)

    @property
    def computing_power(self):
        """Relative speed of this node reported to the master (uniform on an 8xB200 box)."""
        return 0

    # -- introspection -------------------------------------------------------------------------------------
    def print_stats(self, by_name=False, top_number=5, file=None):
        """Top-N unit run times, sum of unit time vs wall clock, per-method totals."""
        timers = {}
        for u in self._units:
            key = u.name if by_name else u
            timers[key] = timers.get(key, 0.0) + u.total_run_time
        stats = sorted(timers.items(), key=lambda kv: kv[1], reverse=True)
        total = sum(t for _, t in stats)
        wall = self._run_time_ if self._run_time_ > 0 else 1e-12
        lines = ["Unit run time statistics top (%d of %d):" % (min(top_number, len(stats)), len(stats))]
        for key, t in stats[:top_number]:
            name = key if by_name else key.name
            lines.append("  %-32s %10.4f s  %5.1f %%" % (name, t, 100.0 * t / total if total else 0.0))
        lines.append("  %-32s %10.4f s  (wall %.4f s, efficiency %.1f %%)"
                     % ("Total", total, wall, 100.0 * total / wall))
        for m, t in sorted(self._method_time_.items()):
            lines.append("  %-32s %10.4f s" % (m, t))
        text = "\n".join(lines)
        if file is not None:
            print(text, file=file)
        else:
            self.info("%s", text)
        return stats

    def generate_graph(self, filename=None, write_on_disk=True, with_data_links=False, background="transparent"):
        """Graphviz DOT text of the control-flow graph (optionally attribute links)."""
        order = []
        seen = set()
        for u in list(self.start_point.dependent_units()) + self._units:
            if id(u) not in seen:
                seen.add(id(u))
                order.append(u)
        ids = {id(u): "n%d" % i for i, u in enumerate(order)}
        out = ["digraph Workflow {", '    bgcolor="%s";' % background, '    mindist="0.5";',
               '    outputorder="edgesfirst";', '    overlap="false";']
        for u in order:
            color = self.VIEW_GROUP_COLORS.get(u.view_group, "white")
            shape = "rect"
            out.append('    %s [label="%s\\n%s", style="filled", fillcolor="%s", shape="%s"];'
                       % (ids[id(u)], u.name, type(u).__name__, color, shape))
        for u in order:
            for dst in u.links_to_sorted:
                if id(dst) not in ids:
                    continue
                attrs = []
                if dst.gate_block._op is not None or bool(dst.gate_block):
                    attrs.append('color="red"')
                out.append("    %s -> %s%s;" % (ids[id(u)], ids[id(dst)],
                                                (" [" + ", ".join(attrs) + "]") if attrs else ""))
        if with_data_links:
            for u in order:
                for k, v in u.__dict__.items():
                    if k.startswith("__link_") and id(v[0]) in ids:
                        out.append('    %s -> %s [style="dashed", label="%s", fontsize=8];'
                                   % (ids[id(v[0])], ids[id(u)], k[len("__link_"):]))
        out.append("}")
        text = "\n".join(out) + "\n"
        if write_on_disk and filename:
            with open(filename, "w") as f:
                f.write(text)
            self.info("Saved the workflow graph to %s", filename)
        return text

    def generate_graph_svg(self, box=(150, 34), gap=(26, 30)):
        """The control-flow graph as a self-contained SVG picture (the reference's web status lays the DOT text out in
        the browser with viz.js; here the layout is done once, on this side).  Layers = longest path from the start
        point along forward edges (an edge into an already placed ancestor - the Repeater loop - is a back edge, drawn
        dashed); inside a layer units are ordered by the mean position of their parents.  Gate-blocked edges are red."""
        import html
        order, seen = [], set()
        for u in list(self.start_point.dependent_units()) + self._units:
            if id(u) not in seen and u is not self:
                seen.add(id(u))
                order.append(u)
        index = {id(u): i for i, u in enumerate(order)}
        edges = [(index[id(u)], index[id(d)], d) for u in order for d in u.links_to_sorted if id(d) in index]
        # depth-first classification of back edges, then longest-path layering over the remaining DAG
        state, back = {}, set()
        children = {}
        for a, b, _ in edges:
            children.setdefault(a, []).append(b)
        for root_i in range(len(order)):
            if root_i in state:
                continue
            stack = [(root_i, iter(children.get(root_i, ())))]
            state[root_i] = 1
            while stack:
                node, it = stack[-1]
                nxt = next(it, None)
                if nxt is None:
                    state[node] = 2
                    stack.pop()
                elif state.get(nxt) == 1:
                    back.add((node, nxt))
                elif nxt not in state:
                    state[nxt] = 1
                    stack.append((nxt, iter(children.get(nxt, ()))))
        layer = [0] * len(order)
        for _ in range(len(order)):
            changed = False
            for a, b, _d in edges:
                if (a, b) not in back and layer[b] < layer[a] + 1:
                    layer[b] = layer[a] + 1
                    changed = True
            if not changed:
                break
        rows = {}
        for i, l in enumerate(layer):
            rows.setdefault(l, []).append(i)
        pos = {}
        for l in sorted(rows):
            members = rows[l]
            if l > 0:
                def bary(i):
                    ps = [pos[a][0] for a, b, _d in edges if b == i and a in pos and (a, b) not in back]
                    return sum(ps) / len(ps) if ps else 0.0
                members.sort(key=bary)
            for k, i in enumerate(members):
                pos[i] = (k, l)
        width = max(len(m) for m in rows.values())
        bw, bh = box
        gx, gy = gap
        W = width * (bw + gx) + gx
        H = (max(rows) + 1) * (bh + gy) + gy

        def xy(i):
            k, l = pos[i]
            pad = (width - len(rows[l])) * (bw + gx) / 2.0
            return gx + pad + k * (bw + gx), gy + l * (bh + gy)

        out = ['<svg xmlns="http://www.w3.org/2000/svg" width="%d" height="%d" viewBox="0 0 %d %d" '
               'font-family="sans-serif" font-size="10">' % (W, H, W, H),
               '<defs><marker id="arr" markerWidth="8" markerHeight="8" refX="7" refY="3" orient="auto">'
               '<path d="M0,0 L7,3 L0,6 z" fill="context-stroke"/></marker></defs>']
        for a, b, dst in edges:
            xa, ya = xy(a)
            xb, yb = xy(b)
            blocked = dst.gate_block._op is not None or bool(dst.gate_block)
            color = "#c00" if blocked else "#555"
            if (a, b) in back:      # around the right-hand side, dashed
                x0, y0, x1, y1 = xa + bw, ya + bh / 2.0, xb + bw, yb + bh / 2.0
                bend = max(x0, x1) + gx * 0.8
                out.append('<path d="M%.1f,%.1f C%.1f,%.1f %.1f,%.1f %.1f,%.1f" fill="none" stroke="%s" '
                           'stroke-dasharray="4,3" marker-end="url(#arr)"/>' % (x0, y0, bend, y0, bend, y1, x1, y1, color))
            else:
                x0, y0, x1, y1 = xa + bw / 2.0, ya + bh, xb + bw / 2.0, yb
                out.append('<path d="M%.1f,%.1f C%.1f,%.1f %.1f,%.1f %.1f,%.1f" fill="none" stroke="%s" '
                           'marker-end="url(#arr)"/>' % (x0, y0, x0, y0 + gy / 2.0, x1, y1 - gy / 2.0, x1, y1, color))
        for i, u in enumerate(order):
            x, y = xy(i)
            fill = self.VIEW_GROUP_COLORS.get(u.view_group, "white")
            out.append('<g><title>%s</title><rect x="%.1f" y="%.1f" width="%d" height="%d" rx="4" fill="%s" '
                       'stroke="#333"/><text x="%.1f" y="%.1f" text-anchor="middle">%s</text>'
                       '<text x="%.1f" y="%.1f" text-anchor="middle" fill="#444" font-size="8">%s</text></g>'
                       % (html.escape("%s (%s)" % (u.name, type(u).__name__)), x, y, bw, bh, fill, x + bw / 2.0, y + 14,
                          html.escape(str(u.name)[:26]), x + bw / 2.0, y + 26, html.escape(type(u).__name__[:30])))
        out.append("</svg>")
        return "\n".join(out)

    def gather_results(self):
        results = {"id": getattr(self.launcher, "id", None), "log_id": getattr(self.launcher, "log_id", None)}
        methods = interface_methods(IResultProvider)
        for u in self._units:
            if all(callable(getattr(u, m, None)) for m in methods):
                names = set(u.get_metric_names())
                values = u.get_metric_values()
                if isinstance(values, dict):
                    missing = names - set(values)
                    if missing:
                        self.warning("%s promised metrics %s but did not deliver them", u.name, sorted(missing))
                    results.update(values)
        return results

    def write_results(self, file=None):
        from .json_encoders import NumpyJSONEncoder
        file = file or self._result_file
        results = self.gather_results()
        if hasattr(file, "write"):
            json.dump(results, file, indent=4, sort_keys=True, cls=NumpyJSONEncoder)
            return results
        os.makedirs(os.path.dirname(os.path.abspath(file)) or ".", exist_ok=True)
        with open(file, "w") as f:
            json.dump(results, f, indent=4, sort_keys=True, cls=NumpyJSONEncoder)
        self.info("Wrote results to %s", file)
        return results

    @property
    def checksum(self):
        """sha1 of the defining source file + "_<number of units>" (handshake compatibility check)."""
        try:
            src = inspect.getsourcefile(type(self))
            with open(src, "rb") as f:
                digest = hashlib.sha1(f.read()).hexdigest()
        except (TypeError, OSError):
            digest = hashlib.sha1(type(self).__name__.encode()).hexdigest()
        return "%s_%d" % (digest, len(self._units))

    # -- export for the native inference runtime ----------------------------------------------------------
    def package_export(self, file_name, archive_format="zip", precision=32):
        """Write ``contents.json`` + ``NNNN_shape.npy`` arrays into a zip / tgz package that the
        C++ runtime (``native/``) loads.  Format is byte-compatible with the reference
        (``veles/workflow.py:868-975``; sample ``libVeles/tests/workflow_files/contents.json``)."""
        if archive_format not in ("zip", "tgz"):
            raise ValueError('Only "zip" and "tgz" formats are supported (got %s)' % archive_format)
        if precision not in (16, 32):
            raise ValueError("Only 16-bit and 32-bit floats are supported (got %s)" % precision)
        exported = [u for u in self.units_in_dependency_order if hasattr(u, "package_export") and u is not self]
        if not exported:
            raise ValueError("No units support export. Implement package_export() in at least one.")
        obj = {"workflow": self.name, "checksum": self.checksum,
               "units": [{"class": {"name": type(u).__name__, "uuid": type(u).__id__},
                          "data": u.package_export()} for u in exported]}
        for ind, unit in enumerate(exported):
            # connect through non-exported units (e.g. dropout) to the next exported ones
            links = []
            seen = set()
            walk = list(unit.links_to_sorted)
            while walk:
                d = walk.pop(0)
                if id(d) in seen:
                    continue
                seen.add(id(d))
                if d in exported:
                    j = exported.index(d)
                    if j > ind and j not in links:
                        links.append(j)
                else:
                    walk.extend(d.links_to_sorted)
            obj["units"][ind]["links"] = sorted(links)
        for i, u in enumerate(obj["units"][:-1]):
            if not u["links"]:
                raise VelesException("Unit %s is not connected to any other unit" % exported[i])
        arrays = []

        def array_name(arr, index, json_mode):
            name = "%04d_%s" % (index, "x".join(map(str, arr.shape)))
            return "@" + name if json_mode else name + ".npy"

        def default(o):
            if isinstance(o, numpy.ndarray):
                arrays.append(o)
                return array_name(o, len(arrays) - 1, True)
            if isinstance(o, (numpy.integer,)):
                return int(o)
            if isinstance(o, (numpy.floating,)):
                return float(o)
            raise TypeError("Objects of class %s are not supported" % type(o))

        main = json.dumps(obj, indent=4, sort_keys=True, default=default).encode("utf-8")

        def npy_bytes(arr):
            bio = io.BytesIO()
            numpy.save(bio, numpy.ascontiguousarray(arr).astype("float%d" % precision))
            return bio.getvalue()

        if archive_format == "zip":
            with zipfile.ZipFile(file_name, mode="w", compression=zipfile.ZIP_DEFLATED) as azip:
                azip.writestr("contents.json", main)
                for ind, arr in enumerate(arrays):
                    azip.writestr(array_name(arr, ind, False), npy_bytes(arr))
        else:
            with tarfile.open(file_name, "w:gz") as tar:
                def add(name, payload):
                    ti = tarfile.TarInfo(name)
                    ti.size = len(payload)
                    ti.mode = 0o666
                    tar.addfile(ti, fileobj=io.BytesIO(payload))
                add("contents.json", main)
                for ind, arr in enumerate(arrays):
                    add(array_name(arr, ind, False), npy_bytes(arr))
        self.info("Successfully exported package to %s", file_name)
        return file_name

    def change_unit(self, old, new, save_gates=True):
        """Hot-swap ``old`` with ``new`` keeping control links (and optionally gates)."""
        if isinstance(old, str):
            old = self[old]
        if old is new:
            return new
        srcs = old.links_from_sorted
        dsts = old.links_to_sorted
        old.unlink_all()
        for s in srcs:
            new.link_from(s)
        for d in dsts:
            d.link_from(new)
        if save_gates:
            new.gate_block = old.gate_block
            new.gate_skip = old.gate_skip
            new.ignores_gate = old.ignores_gate
        if old is self.start_point:
            self.start_point = new
        if old is self.end_point:
            self.end_point = new
        self.del_ref(old)
        if new not in self._units:
            new.workflow = self
        return new


class TrivialWorkflow(Workflow):
    hide_from_registry = True
