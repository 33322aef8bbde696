"""Global configuration tree (``root``) with auto-vivifying attribute access.

Behavioural parity with the reference's ``veles/config.py:60-176`` (auto-created child
nodes, ``update`` deep merge where a dict carrying ``"dict": True`` is stored verbatim,
``protect`` read-only keys, ``get``/``validate_kwargs`` helpers) re-implemented around a
``__slots__``-free node class whose children live in one ordinary dict, which makes the
tree trivially picklable / JSON-able and keeps protected-key bookkeeping on the node.
Defaults mirror ``veles/config.py:178-291`` with the engine section re-targeted at B200.
"""
from __future__ import annotations

import os
import pprint
import sys
import tempfile

_INTERNAL = ("_path_", "_nodes_", "_protected_")


class Config(object):
    """One node of the configuration tree."""

    def __init__(self, path):
        object.__setattr__(self, "_path_", path)
        object.__setattr__(self, "_nodes_", {})
        object.__setattr__(self, "_protected_", set())

    # -- tree API ---------------------------------------------------------------------
    @property
    def __path__(self):
        return self._path_

    @property
    def __content__(self):
        return dict(self._nodes_)

    def update(self, value):
        if self is root:
            raise ValueError("Root updates are disabled")
        if isinstance(value, Config):
            value = value.__content__
        if not isinstance(value, dict):
            raise ValueError("Value must be an instance of dict or Config")
        self._merge(value)
        return self

    def _merge(self, tree):
        for k, v in tree.items():
            if isinstance(v, Config):
                getattr(self, k)._merge(v.__content__)
            elif isinstance(v, dict) and not v.get("dict", False):
                node = self._nodes_.get(k)
                if not isinstance(node, Config):
                    node = Config("%s.%s" % (self._path_, k))
                    setattr(self, k, node)
                node._merge(v)
            else:
                if isinstance(v, dict) and "dict" in v:
                    v = {kk: vv for kk, vv in v.items() if kk != "dict"}
                setattr(self, k, v)

    def protect(self, *names):
        """Make the named children read-only."""
        self._protected_.update(names)

    def print_(self, indent=1, width=80, file=sys.stdout):
        print("-" * width, file=file)
        print('Configuration "%s":' % self._path_, file=file)
        pprint.pprint(self.to_dict(), indent=indent, width=width, stream=file)
        print("-" * width, file=file)

    def to_dict(self):
        return {k: (v.to_dict() if isinstance(v, Config) else v) for k, v in self._nodes_.items()}

    # -- attribute protocol -----------------------------------------------------------
    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        nodes = object.__getattribute__(self, "_nodes_")
        if name in ("keys", "values", "items"):
            return getattr(nodes, name)
        try:
            return nodes[name]
        except KeyError:
            child = Config("%s.%s" % (self._path_, name))
            nodes[name] = child
            return child

    def __setattr__(self, name, value):
        if name in _INTERNAL:
            object.__setattr__(self, name, value)
            return
        if name in self._protected_:
            raise AttributeError("Attempted to change the protected configuration setting %s.%s"
                                 % (self._path_, name))
        self._nodes_[name] = value

    def __delattr__(self, name):
        if name in self._protected_:
            raise AttributeError("Attempted to delete the protected configuration setting %s.%s"
                                 % (self._path_, name))
        del self._nodes_[name]

    def __contains__(self, name):
        return name in self._nodes_

    def __iter__(self):
        return iter(self._nodes_)

    def __getitem__(self, item):
        return getattr(self, item)

    def __setitem__(self, item, value):
        setattr(self, item, value)

    def __len__(self):
        return len(self._nodes_)

    def __bool__(self):
        # an empty (auto-created) node is falsy: ``root.x.y or default`` works
        return len(self._nodes_) > 0

    def __repr__(self):
        return '<Config "%s": %r>' % (self._path_, self._nodes_)

    def __getstate__(self):
        return {"path": self._path_, "nodes": self._nodes_, "protected": sorted(self._protected_)}

    def __setstate__(self, state):
        object.__setattr__(self, "_path_", state["path"])
        object.__setattr__(self, "_nodes_", state["nodes"])
        object.__setattr__(self, "_protected_", set(state["protected"]))

    def __eq__(self, other):
        return self is other

    def __hash__(self):
        return id(self)


root = Config("root")


def get(value, default_value=None):
    """Return ``value`` unless it is an undefined (auto-created) config node."""
    if isinstance(value, Config):
        return default_value
    return value


def validate_kwargs(caller, **kwargs):
    for k, v in kwargs.items():
        if isinstance(v, Config) and len(v) == 0:
            caller.warning("Argument '%s' seems to be undefined at %s", k, v.__path__)
            if get(root.common.trace.undefined_configs, False):
                import traceback
                caller.warning("kwargs are: %s", kwargs)
                caller.warning("Stack trace:\n%s", "".join(traceback.format_stack()[:-1]))


__root__ = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
__home__ = os.path.join(os.environ.get("HOME", tempfile.gettempdir()), ".veles_b200")


def _reset_common():
    root.common.update({
        "dirs": {
            "veles": os.path.join(__root__, "veles_b200"),
            "user": __home__,
            "help": os.path.join(__root__, "docs"),
            "cache": os.path.join(__home__, "cache"),
            "snapshots": os.path.join(__home__, "snapshots"),
            "datasets": os.path.join(__home__, "datasets"),
            "logs": os.path.join(__home__, "logs"),
            "results": os.path.join(__home__, "results"),
        },
        "engine": {
            # "b200" (alias "cuda") = hand-written sm_100a kernels; "numpy" = correctness oracle.
            # "ocl" is accepted on the CLI but rejected with a clear error (no OpenCL on B200).
            "backend": os.environ.get("VELES_BACKEND", "auto"),
            # storage/compute precision of activations and GEMM operands on the device:
            # "bf16" (fp32 accumulate, fp32 master weights) is the B200 default;
            # the numpy oracle computes in "float"/"double".
            "precision_type": "bf16",
            "precision_level": 0,
            "test_known_device": False,
            "test_unknown_device": True,
            "test_precision_types": ("float", "double", "bf16"),
            "test_precision_levels": (0, 1),
            "thread_pool": {"minthreads": 2, "maxthreads": 2},
            "network_compression": "gzip",
            "cuda_graphs": True,
            "source_dirs": (os.environ.get("VELES_ENGINE_DIRS", "").split(":") +
                            [os.path.join(__root__, "veles_b200", "csrc")]),
            "device_dirs": ["/usr/share/veles/devices", os.path.join(__home__, "devices"),
                            os.environ.get("VELES_DEVICE_DIRS", "./")],
        },
        "web": {"host": "127.0.0.1", "port": 8090, "log_file": "/tmp/veles_b200_web.log",
                "log_backups": 9, "notification_interval": 1, "pidfile": "/tmp/veles_b200_web.pid",
                "root": os.path.join(__root__, "web"), "drop_time": 30 * 24 * 3600},
        "graphics": {"multicast_address": "239.192.1.1", "blacklisted_ifaces": set(), "matplotlib":
                     {"backend": "Agg"}, "dir": os.path.join(__home__, "plots")},
        "forge": {"service_name": "service", "upload_name": "upload", "fetch_name": "fetch",
                  "manifest": "manifest.json", "server": "http://127.0.0.1:8080/forge"},
        "genetics": {"disable": {"plotting": True, "snapshotting": True, "publishing": True}},
        "ensemble": {"disable": {"plotting": True, "publishing": True}},
        "evaluation_transform": lambda value, name: value,
        "timings": False,
        "trace": {"run": False, "misprints": False, "undefined_configs": False},
        "disable": {"plotting": "unittest" in sys.modules or "pytest" in sys.modules,
                    "snapshotting": False, "publishing": False},
        "exceptions": {"run_after_stop": True},
        "api": {"port": 0, "path": "/api"},
        "snapshot_dir": os.path.join(__home__, "snapshots"),
    })


_reset_common()

# Site overrides, lowest to highest priority (reference: veles/config.py:293-325)
for _site in ("/etc/default/veles_b200", os.path.join(__home__, "site_config.py"),
              os.path.join(os.getcwd(), "site_config.py")):
    if os.path.isfile(_site):
        try:
            _ns = {}
            with open(_site) as _f:
                exec(compile(_f.read(), _site, "exec"), _ns)  # noqa: S102 - trusted local site config
            if callable(_ns.get("update")):
                _ns["update"](root)
        except Exception as _e:  # pragma: no cover - site file problems must not kill imports
            sys.stderr.write("veles_b200: ignoring site config %s: %s\n" % (_site, _e))
