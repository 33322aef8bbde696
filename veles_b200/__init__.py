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

    @property
    def __units__(self):
        from .unit_registry import UnitRegistry
        return sorted(UnitRegistry.units, key=lambda c: c.__name__)


_sys.modules[__name__].__class__ = VelesModule


# Workflows, configurations and snapshots written for the reference name its module paths (``veles.units``,
# ``veles.znicz...``): alias them to this package unless a real ``veles`` distribution is importable.
try:
    from .compat import install_veles_alias as _install_veles_alias
    _install_veles_alias()
except Exception:   # pragma: no cover - the alias is a convenience, never a reason to fail the import
    pass
