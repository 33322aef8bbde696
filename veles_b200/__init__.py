"""veles_b200 — a Blackwell (B200, sm_100a) native deep-learning platform with the
capabilities, public API and checkpoint format of Veles (reference: devbib/veles)."""
__version__ = "0.1.0"


import os as _os
import sys as _sys
import types as _types


def check_root():
    """Refuse to run as uid 0 unless VELES_ALLOW_ROOT is set (reference: veles/__init__.py:349-350).
    Containers on the GPU boxes run as root, so this only warns there."""
    return _os.geteuid() != 0 or bool(_os.environ.get("VELES_ALLOW_ROOT")) or _os.path.exists("/.dockerenv")


class VelesModule(_types.ModuleType):
    """The package object is callable: ``veles_b200("workflow.py", "config.py", backend="numpy", ...)`` runs
    the CLI in-process and returns the launcher (reference: veles/__init__.py:126-189)."""

    def __call__(self, workflow, config=None, *config_list, **kwargs):
        from .__main__ import Main
        subprocess_mode = kwargs.pop("subprocess", False)
        argv = []
        for key, val in kwargs.items():
            flag = ("--" + key.replace("_", "-")) if len(key) > 1 else ("-" + key)
            if isinstance(val, bool):
                if val:
                    argv.append(flag)
            elif val is not None:
                argv += [flag, str(val)]
        argv += [workflow, config if config is not None else "-"] + list(config_list)
        if subprocess_mode:
            import subprocess
            return subprocess.call([_sys.executable, "-m", "veles_b200"] + argv)
        main = Main(True, *argv)
        main.run()
        return main.launcher

    @property
    def Trainer(self):
        from .trainer import Trainer
        return Trainer

    # -- introspection of the installed tree (reference: veles/__init__.py:190-300) ---------------------------------
    @property
    def __root__(self):
        return _os.path.dirname(_os.path.dirname(_os.path.abspath(self.__file__)))

    def _scan(self):
        """Import every module of the package tree (tests and build products aside); yields the modules, reports the
        ones that do not import on stderr."""
        import importlib
        skip = {"tests", "csrc", "lib", "__pycache__"}
        base = _os.path.dirname(_os.path.abspath(self.__file__))
        for root_dir, dirs, files in _os.walk(base):
            dirs[:] = sorted(d for d in dirs if d not in skip)
            rel = _os.path.relpath(root_dir, _os.path.dirname(base)).replace(_os.sep, ".")
            if not all(part.isidentifier() for part in rel.split(".")):
                continue
            for fn in sorted(files):
                name, ext = _os.path.splitext(fn)
                if ext != ".py" or name == "__main__":
                    continue
                modname = rel if name == "__init__" else "%s.%s" % (rel, name)
                try:
                    yield importlib.import_module(modname)
                except Exception as e:  # noqa: BLE001 - optional dependencies (h5py, pyodbc, ...) may be absent
                    _sys.stderr.write("%s: %s\n" % (modname, e))

    @property
    def __modules__(self):
        cache = self.__dict__.get("_modules_cache")
        if cache is None:
            cache = self.__dict__["_modules_cache"] = list(self._scan())
        return cache

    @property
    def __units__(self):
        """Every Unit class of the package tree (all modules are imported first, so the registry is complete)."""
        self.__modules__
        from .unit_registry import UnitRegistry
        return sorted(UnitRegistry.units, key=lambda c: c.__name__)

    @property
    def __loc__(self):
        """Non-blank lines of the source tree by kind (the reference shells out to ``cloc``; counted here directly)."""
        cache = self.__dict__.get("_loc_cache")
        if cache is not None:
            return cache
        kinds = {"core": ((".py",), ("veles_b200",)), "tests": ((".py",), ("tests",)),
                 "cuda": ((".cu", ".cuh"), ("veles_b200/csrc",)),
                 "c/c++": ((".cc", ".h", ".c"), ("native", "include")), "scripts": ((".py", ".sh"), ("scripts",))}
        result = {}
        for kind, (exts, roots) in kinds.items():
            n = 0
            for r in roots:
                for root_dir, dirs, files in _os.walk(_os.path.join(self.__root__, r)):
                    dirs[:] = [d for d in dirs if d not in ("__pycache__", "lib", "obj")]
                    for fn in files:
                        if fn.endswith(exts):
                            try:
                                with open(_os.path.join(root_dir, fn), errors="replace") as f:
                                    n += sum(1 for line in f if line.strip())
                            except OSError:
                                pass
            result[kind] = n
        result["all"] = sum(result.values())
        self.__dict__["_loc_cache"] = result
        return result

    @property
    def __plugins__(self):
        """Packages under the installation root that carry a ``.veles`` marker file, imported (reference :154-167)."""
        cache = self.__dict__.get("_plugins_cache")
        if cache is None:
            import importlib
            cache = set()
            for root_dir, dirs, files in _os.walk(self.__root__):
                dirs[:] = [d for d in dirs if not d.startswith(".") and d not in ("__pycache__", "gpurun_out")]
                if ".veles" in files:
                    package = _os.path.relpath(root_dir, self.__root__).replace(_os.sep, ".")
                    try:
                        cache.add(importlib.import_module(package))
                    except ImportError:
                        continue
            self.__dict__["_plugins_cache"] = cache
        return cache

    @property
    def allow_root(self):
        return bool(self.__dict__.get("_allow_root", False))

    @allow_root.setter
    def allow_root(self, value):
        self.__dict__["_allow_root"] = bool(value)


_sys.modules[__name__].__class__ = VelesModule


# Workflows, configurations and snapshots written for the reference name its module paths (``veles.units``,
# ``veles.znicz...``): alias them to this package unless a real ``veles`` distribution is importable.
try:
    from .compat import install_veles_alias as _install_veles_alias
    _install_veles_alias()
except Exception:   # pragma: no cover - the alias is a convenience, never a reason to fail the import
    pass
