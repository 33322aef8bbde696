"""Unit — node of the dataflow graph.

Parity target: ``veles/units.py:107-926``.  Kept semantics:
  * control-flow links ``link_from`` with *gates*: a unit runs when every incoming link has
    fired since its last run (``open_gate``), unless ``ignores_gate``; ``gate_block`` stops
    propagation into the unit, ``gate_skip`` skips ``run()`` but still propagates (:524-543,
    :782-803);
  * ``run_dependent`` calls a single successor inline and fans several out on the thread
    pool (:485-505);
  * ``initialize`` may return True = "retry me after the others" (:887-898);
  * ``demand(*attrs)`` = must be linked/assigned before initialize (:682-699, :900-913);
  * ``link_attrs(other, "a", ("mine", "yours"), two_way=)`` shares mutable values by
    reference and immutables through ``LinkableAttribute`` (:638-656, :768-780);
  * per-class wall-clock timers (``Unit.timers``), ``run_was_called``, run-after-stop guard;
  * snapshot contract of ``Pickleable`` (``*_`` attributes are not pickled).
New relative to the reference: units expose ``graph_capturable`` so the B200 scheduler can
replay a steady-state chain of device units as one CUDA graph (veles_b200.accelerated_units).
"""
from __future__ import annotations

import collections
import sys
import threading
import time
import uuid
import weakref

from .config import root, get, validate_kwargs
from .distributable import Distributable, TriviallyDistributable
from .mutable import Bool, LinkableAttribute
from .unit_registry import UnitRegistry
from .verified import Verified, Interface


class IUnit(Interface):
    def initialize(self, **kwargs):
        """Prepare for run(); may return True to be re-initialized after the other units."""

    def run(self):
        """Do the job."""


class UnitException(Exception):
    def __init__(self, unit, *args, **kwargs):
        super(UnitException, self).__init__(*args, **kwargs)
        self.unit = unit


class NotInitializedError(UnitException):
    pass


class RunAfterStopError(UnitException):
    pass


_tls = threading.local()


def nothing(*args, **kwargs):
    return None


class _Timers(dict):
    def __missing__(self, key):
        self[key] = 0.0
        return 0.0


_SPIN = {"chars": "|/-\\", "i": 0, "t": 0.0, "enabled": None}


def spin(stream=None):
    """Progress spinner of a terminal run (reference: ``veles/units.py:443-445`` spins after every unit run unless
    ``root.common.disable.spinning_run_progress``): one character redrawn in place, at most ten times a second, only
    when stderr is a terminal (stderr, so that programs whose stdout is data - ``bench.py``'s JSON line, ``--dump-config``
    - stay clean even when run by hand)."""
    st = _SPIN
    if st["enabled"] is None:
        try:
            st["enabled"] = bool(sys.stderr.isatty()) and not get(root.common.disable.spinning_run_progress, False)
        except Exception:  # noqa: BLE001
            st["enabled"] = False
    if not st["enabled"] and stream is None:
        return False
    now = time.perf_counter()
    if now - st["t"] < 0.1 and stream is None:
        return False
    st["t"] = now
    st["i"] = (st["i"] + 1) % len(st["chars"])
    out = stream if stream is not None else sys.stderr
    try:
        out.write(st["chars"][st["i"]] + "\b")
        out.flush()
    except Exception:  # noqa: BLE001 - a closed stdout must not fail a unit
        st["enabled"] = False
        return False
    return True


class Unit(Distributable, Verified, metaclass=UnitRegistry):
    """General unit of a workflow."""

    hide_from_registry = True
    timers = _Timers()
    visible = True
    # device units set this to True when their run() only enqueues stream-ordered GPU work
    graph_capturable = False

    def __init__(self, workflow, **kwargs):
        self._name = kwargs.get("name")
        self._view_group = kwargs.get("view_group", getattr(type(self), "VIEW_GROUP", None))
        self._demanded = set()
        self._id = str(uuid.uuid4())
        self._links_from = {}
        self._links_to = {}
        self._gate_block = Bool(False)
        self._gate_skip = Bool(False)
        self._ignores_gate = Bool(kwargs.get("ignore_gate", False))
        self._run_calls = 0
        self._is_initialized = False
        self._stopped = False
        timings = get(root.common.timings, False)
        if isinstance(timings, (set, frozenset, list, tuple)):
            timings = type(self).__name__ in timings
        self._timings = bool(kwargs.get("timings", timings))
        self._workflow = None
        super(Unit, self).__init__(**kwargs)
        validate_kwargs(self, **kwargs)
        self.verify_interface(IUnit)
        self.workflow = workflow

    def init_unpickled(self):
        super(Unit, self).init_unpickled()
        self._gate_lock_ = threading.Lock()
        self._run_lock_ = threading.Lock()
        self._override_restored_from_snapshot_ = None
        Unit.timers[self._id] = Unit.timers[self._id]
        # instance-level wrappers: timing / bookkeeping happen however run() is reached
        cls = type(self)
        self.__dict__["initialize"] = self._wrap_initialize(cls.initialize.__get__(self))
        self.__dict__["run"] = self._wrap_run(cls.run.__get__(self))
        self.__dict__["stop"] = self._wrap_stop(cls.stop.__get__(self))

    def __getstate__(self):
        state = super(Unit, self).__getstate__()
        for k in ("initialize", "run", "stop"):
            state.pop(k, None)
        # weak references (cycle breakers) cannot be pickled: store strong refs
        state["_links_from"] = {self._deref(u): v for u, v in self._links_from.items()}
        state["_links_to"] = {self._deref(u): v for u, v in self._links_to.items()}
        if self.stripped_pickle:
            state["_links_from"] = {}
            state["_links_to"] = {}
            state["_workflow"] = None
        return state

    def __repr__(self):
        if self._name is not None:
            return '<%s "%s">' % (type(self).__name__, self._name)
        return "<%s at 0x%x>" % (type(self).__name__, id(self))

    def __lt__(self, other):
        # stable ordering for deterministic traversal
        if not isinstance(other, Unit):
            return NotImplemented
        return (self.name, self._id) < (other.name, other._id)

    # -- simple properties ------------------------------------------------------------------
    @property
    def demanded(self):
        return self._demanded

    @property
    def links_from(self):
        return self._links_from

    @property
    def links_to(self):
        return self._links_to

    @property
    def links_from_sorted(self):
        return sorted(self._deref(u) for u in self._links_from)

    @property
    def links_to_sorted(self):
        return sorted(self._deref(u) for u in self._links_to)

    def _gate_prop(name):  # noqa: N805 - helper building the three gate properties
        def getter(self):
            return getattr(self, name)

        def setter(self, value):
            if not isinstance(value, Bool):
                raise TypeError("veles_b200.mutable.Bool type was expected (got %s)" % type(value))
            setattr(self, name, value)
        return property(getter, setter)

    gate_block = _gate_prop("_gate_block")
    gate_skip = _gate_prop("_gate_skip")
    ignores_gate = _gate_prop("_ignores_gate")
    del _gate_prop

    @property
    def workflow(self):
        return self._workflow

    @workflow.setter
    def workflow(self, value):
        if value is None:
            raise ValueError("Unit must have a hosting Workflow")
        if not hasattr(value, "add_ref"):
            raise TypeError("Attempted to set %s as the workflow of %s: it has no add_ref()" % (value, self))
        if self._workflow is not None:
            self._workflow.del_ref(self)
        self._workflow = value
        value.add_ref(self)

    @property
    def launcher(self):
        wf = self._workflow
        while wf is not None and hasattr(wf, "workflow") and hasattr(wf, "start_point"):
            wf = wf.workflow
        return wf

    @property
    def name(self):
        return self._name if self._name is not None else type(self).__name__

    @name.setter
    def name(self, value):
        if value is not None and not isinstance(value, str):
            raise ValueError("Unit name must be a string")
        self._name = value

    @property
    def id(self):
        return self._id

    @property
    def view_group(self):
        return self._view_group

    @view_group.setter
    def view_group(self, value):
        self._view_group = value

    @property
    def timings(self):
        return self._timings

    @property
    def thread_pool(self):
        return self.workflow.thread_pool if self._workflow is not None else None

    @property
    def is_initialized(self):
        return self._is_initialized

    @property
    def is_master(self):
        return bool(getattr(self.workflow, "is_master", False))

    @property
    def is_slave(self):
        return bool(getattr(self.workflow, "is_slave", False))

    @property
    def is_standalone(self):
        return bool(getattr(self.workflow, "is_standalone", True))

    @property
    def interactive(self):
        return bool(getattr(self.workflow, "interactive", False))

    @property
    def testing(self):
        return bool(getattr(self.workflow, "testing", False))

    @property
    def restored_from_snapshot(self):
        ov = self.__dict__.get("_override_restored_from_snapshot_")
        if ov is not None:
            return ov
        if self.__dict__.get("_restored_from_snapshot_") is not None:
            return self._restored_from_snapshot_
        return bool(getattr(self.workflow, "restored_from_snapshot", False)) if self._workflow is not None else False

    @property
    def run_was_called(self):
        return self._run_calls > 0

    @run_was_called.setter
    def run_was_called(self, value):
        if value:
            raise ValueError("You can not set run_was_called to True")
        self._run_calls = 0

    @property
    def stopped(self):
        return self._stopped

    @stopped.setter
    def stopped(self, value):
        self._stopped = bool(value)

    @property
    def total_run_time(self):
        return Unit.timers[self._id]

    @property
    def average_run_time(self):
        return self.total_run_time / self._run_calls if self._run_calls else 0.0

    # -- to be overridden ---------------------------------------------------------------------
    def initialize(self, **kwargs):
        return None

    def run(self):
        return None

    def stop(self):
        return None

    # -- control flow ---------------------------------------------------------------------------
    @staticmethod
    def _deref(u):
        return u() if isinstance(u, weakref.ReferenceType) else u

    @staticmethod
    def _find_key(container, obj):
        for k in container:
            if Unit._deref(k) is obj:
                return k
        return None

    def initialize_dependent(self):
        for unit in self.dependent_units():
            unit.initialize()

    def run_dependent(self):
        """Notify successors: inline when there is one, on the pool when there are several."""
        if self.stopped and not isinstance(self, Container):
            return
        links = self.links_to_sorted
        blocks = [bool(dst.gate_block) for dst in links]
        trace = get(root.common.trace.run, False)
        live = [dst for i, dst in enumerate(links) if not blocks[i] and not dst.gate_block]
        for index, dst in enumerate(live):
            if trace:
                self.debug("%s -> %s (%d/%d) @%s", self, dst, index + 1, len(live),
                           threading.current_thread().name)
            if len(self._links_to) == 1 or self.thread_pool is None:
                self._dispatch(dst)
            else:
                self.thread_pool.callInThread(self._dispatch, dst)

    def _dispatch(self, dst):
        """Trampoline: successors reached from inside a running chain are queued on the
        current thread and executed by the outermost frame, so a Repeater loop of any
        length runs at constant stack depth (the reference recursed: units.py:485-505)."""
        q = getattr(_tls, "queue", None)
        if q is not None:
            q.append((dst, self))
            return
        _tls.queue = q = collections.deque([(dst, self)])
        try:
            while q:
                d, s = q.popleft()
                d._check_gate_and_run(s)
        finally:
            _tls.queue = None

    def dependent_units(self, with_open_gate=False):
        """BFS over control-flow successors, starting with self."""
        yield self
        visited = {id(self)}
        walk = [(child, self) for child in self.links_to_sorted]
        while walk:
            node, parent = walk.pop(0)
            if id(node) in visited or (with_open_gate and not node.open_gate(parent)):
                continue
            yield node
            visited.add(id(node))
            walk.extend((child, node) for child in node.links_to_sorted)

    def open_gate(self, *args):
        """Mark ``args`` as fired; True when the unit may run (all inputs fired → reset)."""
        if self.ignores_gate:
            return True
        with self._gate_lock_:
            if not self._links_from:
                return True
            for src in args:
                key = self._find_key(self._links_from, src)
                if key is not None:
                    self._links_from[key] = True
            if not all(self._links_from.values()):
                return False
            self._close_gate()
        return True

    def _close_gate(self):
        for k in self._links_from:
            self._links_from[k] = False

    def close_gate(self):
        with self._gate_lock_:
            self._close_gate()

    def close_upstream(self):
        for other in self.links_to_sorted:
            key = self._find_key(other._links_from, self)
            if key is not None:
                other._links_from[key] = False
        return self

    def link_from(self, *args):
        """Add control-flow links ``src -> self``.  A back edge closing a cycle is held weakly."""
        with self._gate_lock_:
            for src in args:
                if self._find_key(self._links_from, src) is not None:
                    continue
                self._links_from[src] = False
                src._links_to[self] = False
                if self._reaches(src):
                    # self -> ... -> src already exists, so src -> self closes a loop:
                    # keep this edge weak so reference cycles do not pin the graph
                    del self._links_from[src]
                    del src._links_to[self]
                    self._links_from[weakref.ref(src)] = False
                    src._links_to[weakref.ref(self)] = False
        return self

    def _reaches(self, target):
        """True if ``target`` is reachable from self through strong links_to edges
        (excluding the direct edge target->self just added)."""
        seen = set()
        stack = [self]
        while stack:
            node = stack.pop()
            if id(node) in seen:
                continue
            seen.add(id(node))
            for k in node._links_to:
                if isinstance(k, weakref.ReferenceType):
                    continue
                if k is target:
                    return True
                stack.append(k)
        return False

    def unlink_from(self, *args):
        with self._gate_lock_:
            for src in args:
                k = self._find_key(self._links_from, src)
                if k is not None:
                    del self._links_from[k]
                k = self._find_key(src._links_to, self)
                if k is not None:
                    del src._links_to[k]
        return self

    def unlink_before(self):
        for src in self.links_from_sorted:
            self.unlink_from(src)
        return self

    def unlink_after(self):
        for dst in self.links_to_sorted:
            dst.unlink_from(self)
        return self

    def unlink_all(self):
        self.unlink_before()
        self.unlink_after()
        return self

    def insert_after(self, *chain):
        """Insert self between the last unit of ``chain`` and its former successors."""
        last = chain[-1]
        former = last.links_to_sorted
        for dst in former:
            dst.unlink_from(last)
        self.link_from(*chain)
        for dst in former:
            dst.link_from(self)
        return self

    def derefed_links_to(self):
        return [self._deref(u) for u in self._links_to]

    def derefed_links_from(self):
        return [self._deref(u) for u in self._links_from]

    # -- data links ---------------------------------------------------------------------------------
    @staticmethod
    def is_immutable(value):
        return isinstance(value, (int, float, complex, str, bytes, bool, tuple, frozenset, type(None)))

    def link_attrs(self, other, *args, **kwargs):
        two_way = kwargs.get("two_way", False)
        for arg in args:
            if isinstance(arg, tuple) and len(arg) == 2 and all(isinstance(a, str) for a in arg):
                self._link_attr(other, arg[0], arg[1], two_way)
            elif isinstance(arg, str):
                self._link_attr(other, arg, arg, two_way)
            else:
                raise TypeError(repr(arg) + " is not a valid attributes pair")
        return self

    def _link_attr(self, other, mine, yours, two_way):
        if isinstance(other, Container) and not hasattr(other, yours):
            setattr(other, yours, False)
        try:
            attr = getattr(other, yours)
        except AttributeError:
            self.error("Unable to link %s.%s to %s.%s", other, yours, self, mine)
            raise
        self._demanded.discard(mine)
        if Unit.is_immutable(attr):
            LinkableAttribute.install(self, mine, other, yours, two_way=two_way)
        else:
            LinkableAttribute.uninstall(self, mine)
            setattr(self, mine, attr)

    def demand(self, *args):
        """Declare attributes that must be set (linked) before initialize()."""
        for attr in args:
            if not isinstance(attr, str):
                raise TypeError("Each demanded attribute must be a string (got %r)" % (attr,))
            self._demanded.add(attr)
            if not self._has_value(attr):
                self.__dict__.setdefault(attr, None)

    def undemand(self, *args):
        for attr in args:
            self._demanded.discard(attr)

    def _has_value(self, attr):
        try:
            return getattr(self, attr) is not None
        except AttributeError:
            return False

    def describe(self, file=sys.stdout):
        out = ['Unit: "%s" (%s)' % (self.name, type(self).__name__)]
        out.append("  incoming links: %s" % (", ".join(u.name for u in self.links_from_sorted) or "none"))
        out.append("  outgoing links: %s" % (", ".join(u.name for u in self.links_to_sorted) or "none"))
        out.append("  gates: block=%s skip=%s ignore=%s" % (self.gate_block, self.gate_skip, self.ignores_gate))
        print("\n".join(out), file=file)

    # -- wrappers -------------------------------------------------------------------------------------
    def _check_gate_and_run(self, src):
        if not self.open_gate(src):  # gate has priority over skip
            return
        pool = self.thread_pool
        if pool is not None and pool.failure is not None:
            return
        if not self.gate_skip:
            if not self._run_lock_.acquire(False):
                return  # previous run still in flight: discard the notification
            try:
                if not self._is_initialized:
                    self.error("%s is not initialized", self.name)
                    raise NotInitializedError(self, "%s is not initialized" % self)
                chain = self.__dict__.get("graph_chain_")
                if chain is not None:
                    chain.run_member(self)     # CUDA-graph replay of a steady-state unit sequence (parallel/step.py)
                else:
                    self.run()
            finally:
                self._run_lock_.release()
        self.run_dependent()

    def _wrap_initialize(self, fn):
        def initialize(**kwargs):
            missing = [a for a in sorted(self._demanded) if not self._has_value(a)]
            if missing:
                raise AttributeError("Attribute(s) %s of unit %s are not linked (demanded)"
                                     % (", ".join(missing), self))
            self._stopped = False
            retry = fn(**kwargs)
            if not retry:
                self._is_initialized = True
            return retry
        initialize.__wrapped__ = fn
        return initialize

    def _wrap_run(self, fn):
        def run(*args, **kwargs):
            if self._stopped and not isinstance(self, Container):
                if get(root.common.exceptions.run_after_stop, True):
                    raise RunAfterStopError(self, "%s's run() was called after stop()" % self)
                self.warning("run() was called after stop()")
                return None
            nvtx = _nvtx_push(self)
            t0 = time.perf_counter()
            try:
                return fn(*args, **kwargs)
            finally:
                dt = time.perf_counter() - t0
                if nvtx is not None:
                    nvtx.range_pop()
                Unit.timers[self._id] += dt
                self._run_calls += 1
                spin()
                if self._timings:
                    self.debug("%s run took %.6f s", self.name, dt)
        run.__wrapped__ = fn
        return run

    def _wrap_stop(self, fn):
        def stop(*args, **kwargs):
            self._stopped = True
            return fn(*args, **kwargs)
        stop.__wrapped__ = fn
        return stop


_NVTX = [None]   # None = not probed; False = disabled; else torch.cuda.nvtx


def _nvtx_push(unit):
    """NVTX range per unit run (``root.common.trace.nvtx = True`` or ``VELES_NVTX=1``): profiler timelines
    (nsys / ncu --nvtx) then show the dataflow graph's units over the kernels they launch."""
    mod = _NVTX[0]
    if mod is None:
        import os
        mod = False
        if get(root.common.trace.nvtx, False) or os.environ.get("VELES_NVTX") == "1":
            try:
                import torch
                if torch.cuda.is_available():
                    mod = torch.cuda.nvtx
            except ImportError:
                pass
        _NVTX[0] = mod
    if mod is False:
        return None
    mod.range_push("%s:%s" % (type(unit).__name__, unit.name))
    return mod


class TrivialUnit(Unit, TriviallyDistributable):
    hide_from_registry = True

    def initialize(self, **kwargs):
        pass

    def run(self):
        pass


class Container(Unit):
    """Unit that contains other units (Workflow)."""
    hide_from_registry = True
