"""Lightweight interface verification (reference: veles/verified.py:45-66 used
zope.interface; here an interface is a plain class whose public callables are the
required method names — no third-party dependency)."""
from .logger import Logger


class Interface(object):
    """Marker base for interface declarations."""


def interface_methods(iface):
    names = []
    for klass in iface.__mro__:
        if klass in (Interface, object):
            continue
        for name, val in vars(klass).items():
            if not name.startswith("_") and callable(val):
                names.append(name)
    return sorted(set(names))


class InterfaceError(TypeError):
    pass


class Verified(Logger):
    """Base for objects that can check they implement an interface."""

    DISABLE_INTERFACE_VERIFICATION = False

    def verify_interface(self, iface):
        if getattr(type(self), "DISABLE_INTERFACE_VERIFICATION", False):
            return
        missing = [m for m in interface_methods(iface) if not callable(getattr(self, m, None))]
        if missing:
            raise InterfaceError("%s does not implement %s: missing %s"
                                 % (type(self).__name__, iface.__name__, ", ".join(missing)))

    @staticmethod
    def verify_class(cls, iface):
        missing = [m for m in interface_methods(iface) if not callable(getattr(cls, m, None))]
        if missing:
            raise InterfaceError("%s does not implement %s: missing %s"
                                 % (cls.__name__, iface.__name__, ", ".join(missing)))


def implementer(*ifaces):
    """Class decorator recording implemented interfaces (documentation + verification aid)."""
    def wrap(cls):
        cls.__implemented__ = tuple(getattr(cls, "__implemented__", ())) + tuple(ifaces)
        return cls
    return wrap
