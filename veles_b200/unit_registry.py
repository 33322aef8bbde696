"""Unit catalogue + constructor-kwargs misprint detection.

Parity target: ``veles/unit_registry.py:51-179``.  The reference extracts the accepted
kwargs from ``__init__`` *bytecode* (``dis`` opcode numbers, CPython-version specific);
here they come from ``inspect.signature`` plus an AST scan of ``kwargs.get("x")`` /
``kwargs["x"]`` / ``kwargs.pop("x")`` patterns over the whole MRO.  At construction a
warning is issued for keyword arguments that no class in the MRO consumes, with a
"did you mean" hint when the Damerau-Levenshtein distance to a known name is 1.
"""
from __future__ import annotations

import ast
import inspect
import textwrap
import uuid


def damerau_levenshtein(a, b):
    """Optimal-string-alignment distance (insert, delete, substitute, transpose)."""
    la, lb = len(a), len(b)
    prev2 = None
    prev = list(range(lb + 1))
    for i in range(1, la + 1):
        cur = [i] + [0] * lb
        for j in range(1, lb + 1):
            cost = 0 if a[i - 1] == b[j - 1] else 1
            cur[j] = min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + cost)
            if i > 1 and j > 1 and a[i - 1] == b[j - 2] and a[i - 2] == b[j - 1]:
                cur[j] = min(cur[j], prev2[j - 2] + 1)
        prev2, prev = prev, cur
    return prev[lb]


def _kwargs_from_init(init):
    """Names a single ``__init__`` accepts: explicit parameters + keys read from **kwargs."""
    names = set()
    try:
        sig = inspect.signature(init)
    except (TypeError, ValueError):
        return names, None
    kwname = None
    for p in sig.parameters.values():
        if p.kind in (p.POSITIONAL_OR_KEYWORD, p.KEYWORD_ONLY) and p.name not in ("self", "workflow"):
            names.add(p.name)
        elif p.kind == p.VAR_KEYWORD:
            kwname = p.name
    if kwname is None:
        return names, None
    try:
        tree = ast.parse(textwrap.dedent(inspect.getsource(init)))
    except (OSError, TypeError, SyntaxError, IndentationError):
        return names, kwname
    for node in ast.walk(tree):
        if isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) \
                and isinstance(node.func.value, ast.Name) and node.func.value.id == kwname \
                and node.func.attr in ("get", "pop", "setdefault") and node.args \
                and isinstance(node.args[0], ast.Constant) and isinstance(node.args[0].value, str):
            names.add(node.args[0].value)
        elif isinstance(node, ast.Subscript) and isinstance(node.value, ast.Name) \
                and node.value.id == kwname:
            sl = node.slice
            if isinstance(sl, ast.Constant) and isinstance(sl.value, str):
                names.add(sl.value)
        elif isinstance(node, ast.Compare) and len(node.ops) == 1 and isinstance(node.ops[0], ast.In) \
                and isinstance(node.left, ast.Constant) and isinstance(node.left.value, str) \
                and isinstance(node.comparators[0], ast.Name) and node.comparators[0].id == kwname:
            names.add(node.left.value)
    return names, kwname


class UnitRegistry(type):
    """Metaclass recording every Unit subclass and its accepted constructor kwargs."""

    units = set()
    hidden_units = set()
    by_uuid = {}
    enabled = True

    def __init__(cls, name, bases, clsdict):
        if not clsdict.get("hide_from_registry", False) and UnitRegistry.enabled:
            UnitRegistry.units.add(cls)
        else:
            UnitRegistry.hidden_units.add(cls)
        if "__id__" not in clsdict:
            # stable id derived from the import path: used by package_export / the C++ runtime
            cls.__id__ = str(uuid.uuid5(uuid.NAMESPACE_DNS, "%s.%s" % (clsdict.get("__module__"), name)))
        UnitRegistry.by_uuid[cls.__id__] = cls
        kwattrs = set()
        init = clsdict.get("__init__")
        if init is not None:
            kwattrs, _ = _kwargs_from_init(init)
        for base in cls.__mro__[1:]:
            kwattrs |= getattr(base, "KWATTRS", set())
        kwattrs |= set(clsdict.get("EXTRA_KWATTRS", ()))
        cls.KWATTRS = kwattrs
        super(UnitRegistry, cls).__init__(name, bases, clsdict)

    def __call__(cls, *args, **kwargs):
        obj = super(UnitRegistry, cls).__call__(*args, **kwargs)
        UnitRegistry._check_misprints(cls, obj, kwargs)
        return obj

    @staticmethod
    def _check_misprints(cls, obj, kwargs):
        known = cls.KWATTRS
        ignored = [k for k in kwargs if k not in known]
        if not ignored:
            return
        warn = getattr(obj, "warning", None)
        if warn is None:
            return
        for k in ignored:
            hints = [n for n in known if damerau_levenshtein(k, n) == 1]
            if hints:
                warn("creating %s: potential misprint in keyword argument name: expected %s - got %s",
                     cls.__name__, " or ".join(sorted(hints)), k)
            else:
                warn("creating %s: unknown keyword argument %s is ignored", cls.__name__, k)


class MappedUnitRegistry(UnitRegistry):
    """UnitRegistry + ``MAPPING`` name registry (see mapped_object_registry)."""
    mapping = "You must define \"mapping\" static attribute in your metaclass"
    base = object

    def __init__(cls, name, bases, clsdict):
        super(MappedUnitRegistry, cls).__init__(name, bases, clsdict)
        mapping = clsdict.get("MAPPING", None)
        meta = type(cls)
        registry = meta.__dict__.get(meta.mapping)
        if registry is None:
            registry = {}
            setattr(meta, meta.mapping, registry)
        if mapping is not None:
            registry[mapping] = cls
