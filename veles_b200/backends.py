"""Device registry: ``b200`` (hand-written sm_100a kernels) and ``numpy`` (oracle).

Parity target: ``veles/backends.py:63-948`` — ``Device()`` is a factory that picks the best
available backend by ``PRIORITY`` unless ``-a/--backend`` / ``VELES_BACKEND`` /
``root.common.engine.backend`` names one; ``-d/--device`` selects the ordinal;
``assign_backend_methods(obj, names)`` binds ``<backend>_<name>`` to ``_backend_<name>_``.
The OpenCL backend and its GEMM auto-tuner (``:623-731``) have no B200 equivalent: ``ocl`` is
recognised and rejected with a clear error.  ``cuda`` is an alias of ``b200``.  There is no
per-thread cuBLAS handle or context push/pop (``:811-836``): torch owns the context and all
kernels are stream-ordered launches of ``lib/libveles_b200.so``.
"""
from __future__ import annotations

import argparse
import os
import threading

from .cmdline import CommandLineArgumentsRegistry
from .config import root, get
from .error import BackendNotSupportedError, DeviceNotFoundError
from .logger import Logger


class BackendRegistry(CommandLineArgumentsRegistry):
    backends = {}

    def __init__(cls, name, bases, clsdict):
        super(BackendRegistry, cls).__init__(name, bases, clsdict)
        backend = clsdict.get("BACKEND")
        if backend is not None:
            BackendRegistry.backends[backend] = cls
            for alias in clsdict.get("ALIASES", ()):
                BackendRegistry.backends[alias] = cls

    @staticmethod
    def backends_as_str():
        return ", ".join(sorted(BackendRegistry.backends))


class DeviceInfo(object):
    """Static description of a device (the reference kept tuned OpenCL block sizes here)."""

    def __init__(self, **kwargs):
        self.desc = kwargs.get("desc", "")
        self.memsize = kwargs.get("memsize", 0)
        self.sm_count = kwargs.get("sm_count", 0)
        self.cc = kwargs.get("cc", (0, 0))

    def __repr__(self):
        return "<DeviceInfo %s, %d SMs, %.1f GB, cc %d.%d>" % (
            self.desc, self.sm_count, self.memsize / 2.0 ** 30, self.cc[0], self.cc[1])


class Device(Logger, metaclass=BackendRegistry):
    """Base device; ``Device(...)`` dispatches to the selected backend class."""
    BACKEND = None
    PRIORITY = 0

    def __new__(cls, *args, **kwargs):
        if cls is not Device:
            return super(Device, cls).__new__(cls)
        backend = kwargs.get("backend") or os.environ.get("VELES_BACKEND") or \
            get(root.common.engine.backend, "auto")
        klass = BackendRegistry.backends.get(backend)
        if klass is None:
            raise ValueError("Unknown backend \"%s\"; available: %s" % (backend, BackendRegistry.backends_as_str()))
        if klass is AutoDevice:
            klass = AutoDevice.choose()
        return super(Device, cls).__new__(klass)

    def __init__(self, **kwargs):
        super(Device, self).__init__()
        self._pid_ = os.getpid()
        self._temp_buffer_size = 0
        self.device_info = DeviceInfo()

    @staticmethod
    def init_parser(**kwargs):
        parser = kwargs.get("parser", argparse.ArgumentParser())
        parser.add_argument("-a", "--backend", type=str, default=os.environ.get("VELES_BACKEND", "auto"),
                            help="Acceleration backend: auto, b200 (alias cuda), numpy.")
        parser.add_argument("-d", "--device", type=str, default="",
                            help="Device ordinal (\"0\", or \"0-3\" ranges when spawning ranks).")
        return parser

    @staticmethod
    def iterparse(spec):
        """Expand "cuda:0-3x2,5" -> ["0","0","1","1","2","2","3","3","5"] (reference :299-308)."""
        out = []
        if ":" in spec:
            spec = spec.split(":", 1)[1]
        for part in filter(None, spec.split(",")):
            mul = 1
            if "x" in part:
                part, m = part.split("x")
                mul = int(m)
            if "-" in part:
                lo, hi = part.split("-")
                ids = range(int(lo), int(hi) + 1)
            else:
                ids = [int(part)]
            for i in ids:
                out.extend([str(i)] * mul)
        return out

    @property
    def backend_name(self):
        return type(self).BACKEND

    @property
    def id(self):
        return None

    @property
    def exists(self):
        return False

    @property
    def is_async(self):
        return False

    @property
    def pid(self):
        return self._pid_

    def sync(self):
        pass

    def thread_pool_attach(self, pool):
        pass

    def thread_pool_detach(self, pool):
        pass

    def request_temp_buffer(self, size):
        self._temp_buffer_size = max(self._temp_buffer_size, size)

    def method_prefixes(self):
        return (self.BACKEND,)

    def assign_backend_methods(self, obj, backend_methods):
        """Bind ``obj.<backend>_<m>`` to ``obj._backend_<m>_`` for every name in ``backend_methods``."""
        for suffix in backend_methods:
            chosen = None
            for prefix in self.method_prefixes():
                chosen = getattr(obj, prefix + "_" + suffix, None)
                if chosen is not None:
                    break
            if chosen is None:
                raise AttributeError("No implementation of \"%s\" for backend %s in %s"
                                     % (suffix, self.backend_name, type(obj).__name__))
            setattr(obj, "_backend_" + suffix + "_", chosen)

    def __getstate__(self):
        return {"backend": self.BACKEND}

    def __setstate__(self, state):
        self.__init__()


class AutoDevice(Device):
    BACKEND = "auto"
    PRIORITY = 0

    @staticmethod
    def available():
        return False

    @staticmethod
    def choose():
        best = None
        for klass in set(BackendRegistry.backends.values()):
            if klass in (Device, AutoDevice) or not klass.available():
                continue
            if best is None or klass.PRIORITY > best.PRIORITY:
                best = klass
        return best or NumpyDevice


class NumpyDevice(Device):
    """Host execution: the correctness oracle (and the CPU tier of the tests)."""
    BACKEND = "numpy"
    PRIORITY = 10

    def __init__(self, **kwargs):
        super(NumpyDevice, self).__init__(**kwargs)
        self.device_info = DeviceInfo(desc="numpy")

    @staticmethod
    def available():
        return True


class OpenCLDevice(Device):
    """Recognised only to produce a clear error: there is no OpenCL path on B200."""
    BACKEND = "ocl"
    PRIORITY = -1

    def __init__(self, **kwargs):
        raise BackendNotSupportedError(
            "The OpenCL backend of the reference has no equivalent here: use -a b200 (alias cuda) or -a numpy")

    @staticmethod
    def available():
        return False


class B200Device(Device):
    """One Blackwell GPU driven through torch streams + lib/libveles_b200.so."""
    BACKEND = "b200"
    ALIASES = ("cuda",)
    PRIORITY = 30
    _lock = threading.Lock()

    def __init__(self, **kwargs):
        super(B200Device, self).__init__(**kwargs)
        import torch
        if not torch.cuda.is_available():
            raise DeviceNotFoundError("No CUDA device is visible to this process")
        index = kwargs.get("index")
        if index is None:
            spec = str(kwargs.get("device", "") or "")
            index = int(self.iterparse(spec)[0]) if spec else int(os.environ.get("LOCAL_RANK", 0))
        self.index = int(index) % torch.cuda.device_count()
        self.torch_device = torch.device("cuda", self.index)
        torch.cuda.set_device(self.torch_device)
        props = torch.cuda.get_device_properties(self.torch_device)
        self.device_info = DeviceInfo(desc=props.name, memsize=props.total_memory,
                                      sm_count=props.multi_processor_count, cc=(props.major, props.minor))
        if props.major != 10:
            raise DeviceNotFoundError(
                "veles_b200 kernels are compiled for sm_100a only; device %d is %s (cc %d.%d)"
                % (self.index, props.name, props.major, props.minor))
        from .ops import _lib
        _lib.lib()  # fail loudly now if the kernel library is missing
        from .ops import kernels
        kernels.bind_device(self.index)
        precision = get(root.common.engine.precision_type, "bf16")
        self.compute_dtype = torch.bfloat16 if precision in ("bf16", "bfloat16") else torch.float32
        self.info("Using %s", self.device_info)

    @staticmethod
    def available():
        try:
            import torch
            return torch.cuda.is_available() and torch.cuda.get_device_capability(0)[0] == 10
        except Exception:  # noqa: BLE001
            return False

    def method_prefixes(self):
        return ("b200", "cuda")

    @property
    def id(self):
        return self.index

    @property
    def exists(self):
        return True

    @property
    def is_async(self):
        return True

    def sync(self):
        import torch
        torch.cuda.synchronize(self.torch_device)

    @property
    def memory_allocated(self):
        import torch
        return torch.cuda.memory_allocated(self.torch_device)

    @property
    def peak_memory_allocated(self):
        import torch
        return torch.cuda.max_memory_allocated(self.torch_device)


CUDADevice = B200Device
