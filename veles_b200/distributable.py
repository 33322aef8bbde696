"""Snapshot (pickle) contract and the job-level distribution interface.

Parity targets (``veles/distributable.py``):
  * ``Pickleable`` (:48-134): attributes whose name ends with ``_`` are never pickled;
    ``init_unpickled()`` recreates them after ``__init__`` *and* after unpickling;
    ``stripped_pickle`` mode additionally dereferences linked attributes.
  * ``Distributable`` (:136-220): thread-safe wrappers around the data exchange methods with
    a 4-second "possible deadlock" warning, ``has_data_for_slave``, ``negotiates_on_connect``.
  * ``IDistributable`` (:222-281) and ``TriviallyDistributable`` (:284-302).

In the B200 design gradient/weight exchange is a fused device kernel (veles_b200.parallel),
so these methods carry *job-level* traffic only: genetics chromosomes, ensemble members,
loader bookkeeping for the ZeroMQ master/slave control plane.
"""
from __future__ import annotations

import threading

from .logger import Logger
from .verified import Interface


class Pickleable(Logger):
    """Base of everything that goes into a snapshot."""

    def __init__(self, **kwargs):
        self._method_storage_ = {}
        super(Pickleable, self).__init__(**kwargs)
        self.init_unpickled()

    def init_unpickled(self):
        """Recreate runtime-only (``*_``) attributes; call chain must reach here."""
        self.__dict__.setdefault("_method_storage_", {})
        self.stripped_pickle_ = False
        parent = super(Pickleable, self)
        if hasattr(parent, "init_unpickled"):
            parent.init_unpickled()

    @property
    def stripped_pickle(self):
        return self.__dict__.get("stripped_pickle_", False)

    @stripped_pickle.setter
    def stripped_pickle(self, value):
        self.stripped_pickle_ = value

    def __getstate__(self):
        state = {}
        linked_prefix = "__link_"
        for k, v in self.__dict__.items():
            if k.endswith("_") and not k.startswith("__"):
                continue
            if self.stripped_pickle and k.startswith(linked_prefix):
                # dereference: store the current value under the plain name
                target, target_name, _ = v
                state[k[len(linked_prefix):]] = getattr(target, target_name)
                continue
            if callable(v) and getattr(v, "__self__", None) is not None and not isinstance(v, type):
                continue  # bound methods are runtime wiring
            state[k] = v
        return state

    def __setstate__(self, state):
        self.__dict__.update(state)
        self._restore_links()
        self.init_unpickled()

    def _restore_links(self):
        from .mutable import LinkableAttribute
        for k in list(self.__dict__):
            if k.startswith("__link_"):
                name = k[len("__link_"):]
                if not isinstance(type(self).__dict__.get(name), LinkableAttribute):
                    setattr(type(self), name, LinkableAttribute(name))


class IDistributable(Interface):
    """Five-method contract of anything taking part in master/slave job exchange."""

    def generate_data_for_master(self):
        """Slave: data to send to the master after a job."""

    def generate_data_for_slave(self, slave):
        """Master: data for a slave's next job (may raise NoMoreJobs)."""

    def apply_data_from_master(self, data):
        """Slave: consume job data."""

    def apply_data_from_slave(self, data, slave):
        """Master: merge an update."""

    def drop_slave(self, slave):
        """Master: forget everything in flight for a dead slave."""


class Distributable(Pickleable):
    """Adds locking and readiness signalling to the IDistributable methods."""

    DEADLOCK_TIME = 4.0
    negotiates_on_connect = False

    def init_unpickled(self):
        super(Distributable, self).init_unpickled()
        self._data_lock_ = threading.RLock()
        self._data_event_ = threading.Event()
        self._data_event_.set()

    def _data_threadsafe(self, fn, name):
        def wrapped(*args, **kwargs):
            if not self._data_lock_.acquire(timeout=self.DEADLOCK_TIME):
                self.warning("Possible deadlock in %s.%s: waited %.0f s for the data lock",
                             type(self).__name__, name, self.DEADLOCK_TIME)
                self._data_lock_.acquire()
            try:
                return fn(*args, **kwargs)
            finally:
                self._data_lock_.release()
        wrapped.__name__ = name + "_threadsafe"
        return wrapped

    @property
    def has_data_for_slave(self):
        return self._data_event_.is_set()

    @has_data_for_slave.setter
    def has_data_for_slave(self, value):
        if value:
            self._data_event_.set()
        else:
            self._data_event_.clear()

    def wait_for_data_for_slave(self, timeout=None):
        if not self._data_event_.wait(timeout if timeout is not None else self.DEADLOCK_TIME):
            self.warning("Possible deadlock in %s: no data for slave after %.0f s", type(self).__name__,
                         timeout if timeout is not None else self.DEADLOCK_TIME)
            if timeout is None:
                self._data_event_.wait()

    # thread-safe entry points used by Workflow's distributed fan-out
    def generate_data_for_master_threadsafe(self):
        return self._data_threadsafe(self.generate_data_for_master, "generate_data_for_master")()

    def generate_data_for_slave_threadsafe(self, slave):
        return self._data_threadsafe(self.generate_data_for_slave, "generate_data_for_slave")(slave)

    def apply_data_from_master_threadsafe(self, data):
        return self._data_threadsafe(self.apply_data_from_master, "apply_data_from_master")(data)

    def apply_data_from_slave_threadsafe(self, data, slave):
        return self._data_threadsafe(self.apply_data_from_slave, "apply_data_from_slave")(data, slave)

    def drop_slave_threadsafe(self, slave):
        return self._data_threadsafe(self.drop_slave, "drop_slave")(slave)


class TriviallyDistributable(object):
    """No-op IDistributable implementation."""

    def generate_data_for_master(self):
        return None

    def generate_data_for_slave(self, slave):
        return None

    def apply_data_from_master(self, data):
        pass

    def apply_data_from_slave(self, data, slave):
        pass

    def drop_slave(self, slave):
        pass
