"""Mutable lazily-evaluated booleans and linkable attributes.

``Bool`` parity target: ``veles/mutable.py:44-216`` — a mutable boolean whose derived
expressions (``a & b``, ``a | b``, ``a ^ b``, ``~a``) stay live: changing ``a`` later changes
the value of every expression built from it; ``a <<= value|callable`` assigns; ``on_true`` /
``on_false`` callbacks fire on ``touch``.  The reference stores derived expressions as
marshalled code objects (Python-version specific); here an expression is an explicit tree of
``(op, operands)`` nodes, so pickles are version-stable.

``LinkableAttribute`` parity target: ``veles/mutable.py:219-357`` — a class-level data
descriptor that forwards ``obj.name`` to ``other.other_name`` (one- or two-way).
"""
from __future__ import annotations

import weakref

_OPS = {
    "or": lambda a, b: a or b,
    "and": lambda a, b: a and b,
    "xor": lambda a, b: a != b,
}


class Bool(object):
    """A mutable boolean value.

    ``a | b``, ``a & b``, ``a ^ b``, ``~a`` build live derived expressions;
    ``a <<= True|False|Bool|callable`` assigns to a *base* Bool; ``bool(a)``/``int(a)`` evaluate.
    """

    __slots__ = ("_op", "_args", "_value", "_influences", "on_true", "on_false", "__weakref__")

    def __init__(self, value=False):
        self._influences = weakref.WeakSet()
        self.on_true = None
        self.on_false = None
        if isinstance(value, Bool):
            # a live alias of another Bool (identity expression)
            self._op = "id"
            self._args = (value,)
            self._value = None
            value._influences.add(self)
        else:
            self._op = None
            self._args = ()
            self._value = False
            self.__ilshift__(value)

    @classmethod
    def _derived(cls, op, *args):
        res = cls.__new__(cls)
        res._influences = weakref.WeakSet()
        res.on_true = None
        res.on_false = None
        res._op = op
        res._args = args
        res._value = None
        for a in args:
            if isinstance(a, Bool):
                a._influences.add(res)
        return res

    @property
    def is_base(self):
        return self._op is None

    def __bool__(self):
        op = self._op
        if op is None:
            v = self._value
            return bool(v()) if callable(v) else bool(v)
        if op == "id":
            return bool(self._args[0])
        if op == "not":
            return not bool(self._args[0])
        return bool(_OPS[op](bool(self._args[0]), bool(self._args[1])))

    def __int__(self):
        return int(bool(self))

    def __index__(self):
        return int(bool(self))

    def __repr__(self):
        return repr(bool(self))

    __str__ = __repr__

    def __ilshift__(self, value):
        if self._op is not None:
            raise RuntimeError("Derived expressions cannot be assigned to.")
        if isinstance(value, Bool):
            value = bool(value)
        if isinstance(value, (bool, int)) and not callable(value):
            self._value = bool(value)
        elif callable(value):
            self._value = value
        else:
            raise TypeError("Value must be a boolean value or a function")
        self.touch()
        return self

    def __call__(self, other):
        """``a(b)``: copy the current value of ``b`` into ``a``."""
        self.__ilshift__(bool(other))
        return self

    def _binary(self, op, other):
        if not isinstance(other, Bool):
            other = Bool(bool(other))
        return Bool._derived(op, self, other)

    def __or__(self, other):
        return self._binary("or", other)

    def __and__(self, other):
        return self._binary("and", other)

    def __xor__(self, other):
        return self._binary("xor", other)

    __ror__ = __or__
    __rand__ = __and__
    __rxor__ = __xor__

    def __invert__(self):
        return Bool._derived("not", self)

    def touch(self):
        """Fire ``on_true`` / ``on_false`` of this Bool and everything derived from it."""
        seen = set()
        pending = [self]
        order = []
        while pending:
            item = pending.pop()
            if id(item) in seen:
                continue
            seen.add(id(item))
            order.append(item)
            pending.extend(list(item._influences))
        for b in order:
            if b.on_true is None and b.on_false is None:
                continue
            if b:
                if b.on_true is not None:
                    b.on_true(b)
            elif b.on_false is not None:
                b.on_false(b)

    def unref(self, item):
        self._influences.discard(item)

    # callbacks and influence sets are runtime-only; callables are pickled by reference
    def __getstate__(self):
        value = self._value
        if callable(value):
            try:
                import pickle
                pickle.dumps(value)
            except Exception:  # noqa: BLE001 - unpicklable lambda: freeze its current value
                value = bool(value())
        return {"op": self._op, "args": self._args, "value": value}

    def __setstate__(self, state):
        self._influences = weakref.WeakSet()
        self.on_true = None
        self.on_false = None
        self._op = state["op"]
        self._args = tuple(state["args"])
        self._value = state["value"]
        for a in self._args:
            if isinstance(a, Bool):
                a._influences.add(self)


class LinkableAttribute(object):
    """Data descriptor forwarding ``obj.<name>`` to ``<target>.<target_name>``.

    Installed on the *class* of ``obj`` (one descriptor per attribute name); the pointer
    ``(target, target_name, two_way)`` is stored per instance under ``__link_<name>``.
    Instances without a pointer fall back to an ordinary per-instance value.
    """

    def __init__(self, name, assignment_guard=True):
        self.name = name
        self.slot = "__link_" + name
        self.value_slot = name  # unlinked instances keep a plain value under the attribute name
        self.assignment_guard = assignment_guard
        self.has_default = False
        self.default = None

    @staticmethod
    def install(obj, name, target, target_name, two_way=False, assignment_guard=True):
        if obj is target and name == target_name:
            raise ValueError("Attempted to link the attribute %r to itself" % name)
        cls = type(obj)
        desc = cls.__dict__.get(name)
        if not isinstance(desc, LinkableAttribute):
            # keep an existing per-instance value reachable for other instances
            prev = getattr(cls, name, LinkableAttribute)
            desc = LinkableAttribute(name, assignment_guard)
            if prev is not LinkableAttribute and not hasattr(prev, "__get__"):
                desc.has_default, desc.default = True, prev  # plain class-level default
            setattr(cls, name, desc)
        obj.__dict__[desc.slot] = (target, target_name, two_way)
        return desc

    @staticmethod
    def uninstall(obj, name):
        desc = type(obj).__dict__.get(name)
        if isinstance(desc, LinkableAttribute):
            obj.__dict__.pop(desc.slot, None)

    def __get__(self, obj, objtype=None):
        if obj is None:
            return self
        ptr = obj.__dict__.get(self.slot)
        if ptr is None:
            try:
                return obj.__dict__[self.value_slot]
            except KeyError:
                if self.has_default:
                    return self.default
                raise AttributeError("%r object has no attribute %r" % (type(obj).__name__, self.name))
        return getattr(ptr[0], ptr[1])

    def __set__(self, obj, value):
        ptr = obj.__dict__.get(self.slot)
        if ptr is None:
            obj.__dict__[self.value_slot] = value
            return
        target, target_name, two_way = ptr
        if two_way:
            setattr(target, target_name, value)
        elif self.assignment_guard:
            raise RuntimeError("Attempted to set the one-way linked attribute %r of %r; "
                               "link it with two_way=True or unlink first" % (self.name, obj))
        else:
            # break the link, keep a private value
            del obj.__dict__[self.slot]
            obj.__dict__[self.value_slot] = value

    def __delete__(self, obj):
        obj.__dict__.pop(self.slot, None)
        obj.__dict__.pop(self.value_slot, None)


def link(obj, name, target, target_name=None, two_way=False):
    return LinkableAttribute.install(obj, name, target, target_name or name, two_way)
