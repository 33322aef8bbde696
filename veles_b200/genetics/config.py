"""``Range`` placeholders in the config tree and helpers walking it.

Parity target: ``veles/genetics/config.py:45-227``: ``Tuneable(default)``, ``Range(default, min, max)``
put anywhere in ``root`` (including inside lists/dicts such as ``layers``); ``fix_config`` replaces
them by their defaults for normal runs (``veles/__main__.py:719-721``); ``process_config`` visits
them with their address so the optimizer can write gene values back.
"""
from __future__ import annotations

from ..config import Config


class Tuneable(object):
    def __init__(self, default):
        self.default = default
        self.path = None
        self.name = None
        self.addr = None

    @property
    def full_name(self):
        return "%s.%s" % (self.path, self.name)

    def set(self, value):
        container, key = self.addr
        cast = type(self.default)
        container[key] = cast(round(value)) if cast is int else cast(value)

    def __ilshift__(self, value):
        self.set(value)
        return self

    def details(self):
        return "default: %s" % (self.default,)

    def __repr__(self):
        return "%s{%s}" % (type(self).__name__, self.details())


class Range(Tuneable):
    """A tunable numeric range: ``Range(default, min_value, max_value)``."""

    def __init__(self, default, min_value, max_value):
        super(Range, self).__init__(default)
        if not min_value <= default <= max_value:
            raise ValueError("Range default %s is outside [%s, %s]" % (default, min_value, max_value))
        self.min_value = min_value
        self.max_value = max_value

    def set(self, value):
        if value < self.min_value or value > self.max_value:
            raise ValueError("[%s] value %s is out of range [%s, %s]"
                             % (self.full_name, value, self.min_value, self.max_value))
        super(Range, self).set(value)

    def details(self):
        return "[%s, %s] (default: %s)" % (self.min_value, self.max_value, self.default)

    # A config may derive further hyper-parameters from a Range before the genes are fixed
    # (``learning_rate_bias = 2 * lr``): arithmetic with a number gives a Range over the transformed interval.
    def _scaled(self, fn):
        vals = sorted((fn(self.min_value), fn(self.max_value)))
        d = fn(self.default)
        return Range(min(max(d, vals[0]), vals[1]), vals[0], vals[1])

    def __mul__(self, k):
        return self._scaled(lambda v: v * k)

    __rmul__ = __mul__

    def __truediv__(self, k):
        return self._scaled(lambda v: v / k)

    def __add__(self, k):
        return self._scaled(lambda v: v + k)

    __radd__ = __add__

    def __sub__(self, k):
        return self._scaled(lambda v: v - k)

    def __neg__(self):
        return self._scaled(lambda v: -v)

    def __float__(self):
        return float(self.default)


def process_config(config, class_to_process, callback):
    """Apply ``callback(path, (container, key), name, value) -> new value`` to every element of
    ``class_to_process`` found in the Config tree (descending into dicts / lists / tuples)."""
    _walk(config.__path__, config._nodes_, class_to_process, callback)


def _walk(path, items, klass, callback):
    keys = sorted(items, key=str) if isinstance(items, dict) else range(len(items))
    for k in keys:
        v = items[k]
        if isinstance(v, Config):
            _walk(v.__path__, v._nodes_, klass, callback)
        elif isinstance(v, dict):
            _walk("%s.%s" % (path, k), v, klass, callback)
        elif isinstance(v, list):
            _walk("%s.%s" % (path, k), v, klass, callback)
        elif isinstance(v, tuple):
            lst = list(v)
            _walk("%s.%s" % (path, k), lst, klass, callback)
            if any(a is not b for a, b in zip(lst, v)):
                items[k] = tuple(lst)
        elif isinstance(v, klass):
            items[k] = callback(path, (items, k), k, v)


def fix_config(cfgroot):
    """Replace every Tuneable by its default (normal, non-optimizing runs)."""
    process_config(cfgroot, Tuneable, lambda path, addr, name, value: value.default)


def collect_tuneables(cfgroot):
    """List of Tuneables with their path/name/addr filled in (tree is left untouched)."""
    found = []

    def visit(path, addr, name, value):
        value.path, value.name, value.addr = path, str(name), addr
        found.append(value)
        return value
    process_config(cfgroot, Tuneable, visit)
    return found


def print_config(cfgroot):
    for name, cfg in cfgroot.__content__.items():
        if name != "common" and isinstance(cfg, Config):
            cfg.print_()
