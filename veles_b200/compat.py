"""Python-version shims kept for API compatibility with user code written against the reference (veles/compat.py)."""
from_none = lambda exc: exc  # noqa: E731 - `raise from_none(e)` idiom


def is_interactive():
    import sys
    return hasattr(sys, "ps1") or "IPython" in sys.modules


# -- import-path compatibility ---------------------------------------------------------------------------------------
# Workflows and snapshots written against the reference name its modules (``veles.units``, ``veles.znicz.all2all``,
# ``veles.loader.fullbatch`` ...).  ``install_veles_alias()`` makes those names resolve to the corresponding
# ``veles_b200`` modules (same objects, not copies), so ``from veles.znicz.standard_workflow import StandardWorkflow``
# in a user's workflow file and ``pickle.load`` of a snapshot that references ``veles.*`` classes both work
# ("classes must live at the same import paths").  No-op when a real ``veles`` package is importable.
import importlib
import importlib.abc
import importlib.util
import sys as _sys


class _VelesAliasFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    PREFIX = "veles"
    TARGET = "veles_b200"

    def _target(self, fullname):
        if fullname == self.PREFIX:
            return self.TARGET
        if fullname.startswith(self.PREFIX + "."):
            return self.TARGET + fullname[len(self.PREFIX):]
        return None

    def find_spec(self, fullname, path=None, target=None):
        t = self._target(fullname)
        if t is None:
            return None
        try:
            if importlib.util.find_spec(t) is None:
                return None
        except (ImportError, ValueError):
            return None
        return importlib.util.spec_from_loader(fullname, self, is_package=True)

    def create_module(self, spec):
        mod = importlib.import_module(self._target(spec.name))
        return mod

    def exec_module(self, module):
        pass   # the real module is already initialised; the alias shares it


def install_veles_alias(force=False):
    """Make ``import veles[.x.y]`` resolve to ``veles_b200[.x.y]``.  Returns True if the alias is active."""
    if any(isinstance(f, _VelesAliasFinder) for f in _sys.meta_path):
        return True
    if not force:
        try:
            if "veles" in _sys.modules or importlib.util.find_spec("veles") is not None:
                return False   # the real package is present: leave it alone
        except (ImportError, ValueError):
            pass
    _sys.meta_path.insert(0, _VelesAliasFinder())   # ahead of PathFinder: the aliased packages carry real __path__s
    return True
