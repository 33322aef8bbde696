"""Units with a device implementation next to the numpy oracle.

Parity target: ``veles/accelerated_units.py:130-866``: ``AcceleratedUnit`` binds
``<backend>_init`` / ``<backend>_run`` by name when the device is assigned, ``--force-numpy``
and ``--sync-run`` flags, ``init_vectors`` / ``unmap_vectors`` helpers, ``DeviceBenchmark`` and
``AcceleratedWorkflow.computing_power``.

What is gone on purpose: run-time source assembly (``#define`` injection, Jinja2 ``.jcu``
templates), the ``nvcc`` JIT and the tar.gz kernel-binary cache (``:298-419, 509-673``).
Kernels are compiled ahead of time for sm_100a into ``lib/libveles_b200.so`` and selected by
arguments, not by text substitution.  ``build_program`` survives as a no-op that verifies
the library is loadable so user units written against the old API keep working.
"""
from __future__ import annotations

import argparse
import time

import numpy

from .backends import Device, NumpyDevice
from .cmdline import CommandLineArgumentsRegistry
from .config import root, get
from .distributable import TriviallyDistributable
from .memory import Array
from .unit_registry import UnitRegistry
from .units import Unit
from .verified import Interface
from .workflow import Workflow


class INumpyUnit(Interface):
    def numpy_init(self):
        """Prepare the host implementation."""

    def numpy_run(self):
        """Host implementation (the oracle)."""


class IB200Unit(Interface):
    def b200_init(self):
        """Allocate device state."""

    def b200_run(self):
        """Enqueue sm_100a kernels on the current stream."""


ICUDAUnit = IB200Unit


class AcceleratedUnitRegistry(UnitRegistry, CommandLineArgumentsRegistry):
    pass


class AcceleratedUnit(Unit, metaclass=AcceleratedUnitRegistry):
    """Unit that runs on the selected device."""
    hide_from_registry = True
    backend_methods = ("run", "init")
    force_numpy_classes = set()
    sync_run = False

    def __init__(self, workflow, **kwargs):
        self._device = None
        self._force_numpy = kwargs.get("force_numpy", False)
        self.intel_opencl_workaround = False
        super(AcceleratedUnit, self).__init__(workflow, **kwargs)
        self._sync = kwargs.get("sync", AcceleratedUnit.sync_run)

    def init_unpickled(self):
        super(AcceleratedUnit, self).init_unpickled()
        self._backend_run_ = None
        self._backend_init_ = None
        self.device_ = None

    @staticmethod
    def init_parser(**kwargs):
        parser = kwargs.get("parser", argparse.ArgumentParser())
        parser.add_argument("--force-numpy", default="", type=str,
                            help="Comma separated unit class names to run on the numpy oracle.")
        parser.add_argument("--sync-run", default=False, action="store_true",
                            help="Synchronize the device after each unit run (honest per-unit timers).")
        return parser

    @classmethod
    def apply_args(cls, args):
        AcceleratedUnit.force_numpy_classes = set(filter(None, getattr(args, "force_numpy", "").split(",")))
        AcceleratedUnit.sync_run = bool(getattr(args, "sync_run", False))

    @property
    def device(self):
        return self.device_

    @device.setter
    def device(self, value):
        if value is None:
            value = NumpyDevice()
        if not isinstance(value, Device):
            raise TypeError("device must be a veles_b200.backends.Device (got %s)" % type(value))
        if self._force_numpy or type(self).__name__ in AcceleratedUnit.force_numpy_classes:
            value = NumpyDevice()
        self.device_ = value
        value.assign_backend_methods(self, type(self).backend_methods)

    @property
    def force_numpy(self):
        return self._force_numpy

    @property
    def on_device(self):
        return self.device_ is not None and self.device_.exists

    @property
    def sync(self):
        return self._sync

    def initialize(self, device=None, **kwargs):
        self.device = device
        self._backend_init_()
        return None

    def run(self):
        res = self._backend_run_()
        if self._sync and self.device_ is not None:
            self.device_.sync()
        return res

    # oracle defaults so simple units need to define only what they use
    def numpy_init(self):
        pass

    def b200_init(self):
        pass

    # -- helpers kept from the reference API ---------------------------------------------------------
    def init_vectors(self, *vecs):
        for v in vecs:
            if isinstance(v, Array) and v:
                v.initialize(self.device_)

    def unmap_vectors(self, *vecs):
        for v in vecs:
            if isinstance(v, Array):
                v.unmap()

    def build_program(self, *args, **kwargs):
        """Reference API shim: kernels are prebuilt; make sure the library loads."""
        if self.on_device:
            from .ops import _lib
            _lib.lib()

    def get_kernel(self, name):
        from .ops import _lib
        return getattr(_lib.lib(), name)


class TrivialAcceleratedUnit(AcceleratedUnit, TriviallyDistributable):
    hide_from_registry = True

    def numpy_run(self):
        pass

    def b200_run(self):
        pass


class DeviceBenchmark(AcceleratedUnit, TriviallyDistributable):
    """Measures GEMM speed: ``computing power`` = 1000 / (time of `repeats` size^3 GEMMs)
    (reference: ``accelerated_units.py:706-824`` used a 1500^2 cuBLAS sgemm x10)."""
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(DeviceBenchmark, self).__init__(workflow, **kwargs)
        self.size = kwargs.get("size", 1536)
        self.repeats = kwargs.get("repeats", 10)
        self.dry_run_first = kwargs.get("dry_run_first", True)
        self.computing_power = 0.0

    def numpy_run(self):
        a = numpy.ones((self.size, self.size), dtype=numpy.float32)
        b = numpy.ones((self.size, self.size), dtype=numpy.float32)
        if self.dry_run_first:
            a.dot(b)
        t0 = time.perf_counter()
        for _ in range(self.repeats):
            a.dot(b)
        dt = time.perf_counter() - t0
        self.computing_power = 1000.0 / dt
        return dt

    def b200_run(self):
        import torch
        from .ops import kernels
        dev = self.device_.torch_device
        a = torch.ones(self.size, self.size, dtype=torch.bfloat16, device=dev)
        b = torch.ones(self.size, self.size, dtype=torch.bfloat16, device=dev)
        c = torch.empty(self.size, self.size, dtype=torch.bfloat16, device=dev)
        if self.dry_run_first:
            kernels.gemm(a, b, c)
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(self.repeats):
            kernels.gemm(a, b, c)
        e1.record()
        torch.cuda.synchronize(dev)
        dt = e0.elapsed_time(e1) / 1000.0
        self.computing_power = 1000.0 / max(dt, 1e-9)
        return dt


class AcceleratedWorkflow(Workflow):
    """Workflow owning a device; reports ``computing_power`` to the master."""
    hide_from_registry = True
    POWER_INTERVAL = 120.0

    def __init__(self, workflow, **kwargs):
        super(AcceleratedWorkflow, self).__init__(workflow, **kwargs)
        self._power_ = 0
        self._last_power_measurement_time = 0.0
        self.device = None

    def init_unpickled(self):
        super(AcceleratedWorkflow, self).init_unpickled()
        self._power_ = 0
        self._power_measure_time_interval = self.POWER_INTERVAL

    def initialize(self, device=None, **kwargs):
        self.device = device
        return super(AcceleratedWorkflow, self).initialize(device=device, **kwargs)

    @property
    def computing_power(self):
        now = time.time()
        if now - self._last_power_measurement_time > self._power_measure_time_interval or not self._power_:
            self._last_power_measurement_time = now
            bench = DeviceBenchmark(self, name="benchmark")
            try:
                bench.initialize(device=self.device)
                bench.run()
                self._power_ = bench.computing_power
            finally:
                self.del_ref(bench)
        return self._power_
