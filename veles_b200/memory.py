"""Array: a numpy host array paired with a torch device tensor, with the reference's
map/unmap coherence protocol.

Parity target: ``veles/memory.py:110-512`` —
  ``map_read``       device -> host if the device copy is newer
  ``map_write``      same, then mark the host copy dirty
  ``map_invalidate`` mark the host copy dirty without copying (host will be overwritten)
  ``unmap``          host -> device if the host copy is dirty
pickling = ``map_read`` then the plain host array (``:284-299``), so torch tensors never leak
into a snapshot.  ``Watcher`` (``:56-107``) counts device bytes.

B200 specifics: the device tensor may use a narrower dtype than the host array (bf16
activations/weight shadows next to fp32 host values) and 2-D device storage can be *pitched*
(row stride rounded up to 8 elements = 16 bytes) so it can be addressed by TMA descriptors.
"""
from __future__ import annotations

import threading

import numpy

from .distributable import Pickleable



# Per-thread log of the host<->device bookkeeping a traced unit sequence performs (parallel/step.py GraphedChain): a
# CUDA-graph replay repeats the kernels but not this Python-side state, so the chain re-applies it around each replay.
import threading as _threading
SYNC_TRACE = _threading.local()

class Watcher(object):
    """Device memory accounting."""
    _lock = threading.Lock()
    mem_in_use = 0
    max_mem_in_use = 0

    @classmethod
    def add(cls, nbytes):
        with cls._lock:
            cls.mem_in_use += nbytes
            cls.max_mem_in_use = max(cls.max_mem_in_use, cls.mem_in_use)

    @classmethod
    def sub(cls, nbytes):
        with cls._lock:
            cls.mem_in_use -= nbytes

    @classmethod
    def reset(cls):
        with cls._lock:
            cls.mem_in_use = 0
            cls.max_mem_in_use = 0


def roundup(num, align):
    d = num % align
    return num if d == 0 else num + (align - d)


class Array(Pickleable):
    """Host/device array pair."""

    def __init__(self, data=None, shallow_pickle=False, dev_dtype=None, pitched=False, **kwargs):
        self._mem = None
        self.shallow_pickle = shallow_pickle
        self._dev_dtype_name = None
        self._pitched = bool(pitched)
        super(Array, self).__init__(**kwargs)
        if dev_dtype is not None:
            self.dev_dtype = dev_dtype
        if data is not None:
            self.mem = data

    def init_unpickled(self):
        super(Array, self).init_unpickled()
        self.device_ = None
        self.devmem_ = None
        self.devbuf_ = None
        self.host_dirty_ = False   # host newer than device
        self.dev_dirty_ = False    # device newer than host
        self.lock_ = threading.Lock()

    # -- host side --------------------------------------------------------------------------------
    @property
    def mem(self):
        return self._mem

    @mem.setter
    def mem(self, value):
        if value is not None and not isinstance(value, numpy.ndarray):
            value = numpy.asarray(value)
        if self.devmem_ is not None and value is not None and self._mem is not None \
                and value.shape != self._mem.shape:
            raise ValueError("Cannot change the shape of an initialized Array; call reset() first")
        self._mem = value
        if self.devmem_ is not None:
            self.host_dirty_ = True
            self.dev_dirty_ = False

    @property
    def dev_dtype(self):
        if self._dev_dtype_name is None:
            return None
        import torch
        return getattr(torch, self._dev_dtype_name)

    @dev_dtype.setter
    def dev_dtype(self, value):
        self._dev_dtype_name = None if value is None else str(value).replace("torch.", "")

    @property
    def device(self):
        return self.device_

    @property
    def devmem(self):
        # ``lazy_device_`` (optional callable): brings lazily produced device contents up to date before anybody
        # touches the tensor (e.g. pooling's int32 offsets expanded from its 1-byte argmax records)
        hook = getattr(self, "lazy_device_", None)
        if hook is not None:
            self.lazy_device_ = None
            try:
                hook()
            finally:
                self.lazy_device_ = hook
        # pending host writes reach the device before a kernel reads the tensor, whether or not the reading unit
        # remembered its unmap() (the usual producers are device units, so this is one flag test on the hot path)
        if self.host_dirty_ and self.devmem_ is not None:
            self.unmap()
        return self.devmem_

    @property
    def shape(self):
        return self._mem.shape if self._mem is not None else ()

    @property
    def dtype(self):
        return self._mem.dtype if self._mem is not None else None

    @property
    def size(self):
        return self._mem.size if self._mem is not None else 0

    @property
    def nbytes(self):
        return self._mem.nbytes if self._mem is not None else 0

    @property
    def sample_size(self):
        return self.size // self.shape[0] if self._mem is not None and self.shape and self.shape[0] else 0

    @property
    def matrix(self):
        self.map_read()
        return self._mem.reshape(self.shape[0], self.sample_size)

    @property
    def plain(self):
        return self._mem.ravel()

    def __bool__(self):
        return self._mem is not None and self._mem.size > 0

    def __len__(self):
        return self._mem.shape[0] if self._mem is not None else 0

    def __getitem__(self, key):
        self.map_read()
        return self._mem[key]

    def __setitem__(self, key, value):
        self.map_write()
        self._mem[key] = value

    def __repr__(self):
        return "<Array %s %s%s>" % (self.shape, self.dtype, " on %s" % self.devmem_.device if self.devmem_ is not None else "")

    def reset(self, new_mem=None):
        """Drop both copies (optionally install a new host array)."""
        with self.lock_:
            self._free_device()
            self._mem = None
        if new_mem is not None:
            self.mem = new_mem

    def _free_device(self):
        if self.devbuf_ is not None:
            Watcher.sub(self.devbuf_.numel() * self.devbuf_.element_size())
        self.devmem_ = None
        self.devbuf_ = None
        self.host_dirty_ = False
        self.dev_dirty_ = False

    # -- device side ------------------------------------------------------------------------------------
    def initialize(self, device, storage=None):
        """Allocate the device tensor (no-op on the numpy device) and upload the host copy.
        ``storage``: an existing device tensor (e.g. a slice of a parameter arena or of a
        symmetric-memory slab) to adopt instead of allocating."""
        if device is None or not getattr(device, "exists", False):
            self.device_ = device
            return
        if self._mem is None:
            raise ValueError("Array.initialize() needs host data first")
        if self.devmem_ is not None and self.device_ is device and tuple(self.devmem_.shape) == self._mem.shape:
            return
        import torch
        self.device_ = device
        self._free_device()
        dt = self.dev_dtype or _torch_dtype(self._mem.dtype)
        shape = self._mem.shape
        if storage is not None:
            if storage.numel() != self._mem.size:
                raise ValueError("storage has %d elements for an Array of %d" % (storage.numel(), self._mem.size))
            self.devbuf_ = storage
            self.devmem_ = storage.view(shape) if storage.is_contiguous() else storage
            self.host_dirty_ = True
            self.unmap()
            return
        if self._pitched and len(shape) >= 2:
            cols = int(numpy.prod(shape[1:]))
            pitch = roundup(cols, 8)
            self.devbuf_ = torch.zeros(shape[0], pitch, dtype=dt, device=device.torch_device)
            self.devmem_ = self.devbuf_[:, :cols]
        else:
            self.devbuf_ = torch.zeros(shape, dtype=dt, device=device.torch_device)
            self.devmem_ = self.devbuf_
        Watcher.add(self.devbuf_.numel() * self.devbuf_.element_size())
        self.host_dirty_ = True
        self.unmap()

    def _dev_view(self):
        """Device tensor shaped like the host array (2-D for pitched storage)."""
        if self._pitched and self._mem.ndim >= 2:
            return self.devmem_
        return self.devmem_

    def map_read(self):
        hook = getattr(self, "lazy_device_", None)
        if hook is not None:
            hook()
        if self.devmem_ is None or not self.dev_dirty_:
            return
        with self.lock_:
            import torch
            src = self.devmem_
            host = src.detach().to(torch.float32 if src.dtype == torch.bfloat16 else src.dtype).cpu().numpy()
            if self._mem.dtype.kind == "u" and self._mem.dtype.itemsize > 1:
                host = host.view(self._mem.dtype)
            self._mem[...] = host.reshape(self._mem.shape).astype(self._mem.dtype, copy=False)
            self.dev_dirty_ = False

    def map_write(self):
        self.map_read()
        if self.devmem_ is not None:
            self.host_dirty_ = True

    def map_invalidate(self):
        if self.devmem_ is not None:
            self.host_dirty_ = True
            self.dev_dirty_ = False

    def unmap(self):
        if self.devmem_ is None or not self.host_dirty_:
            return
        log = getattr(SYNC_TRACE, "log", None)
        if log is not None:
            log["unmaps"].append(self)       # a real host -> device copy happened inside a traced unit sequence
        with self.lock_:
            import torch
            hm = numpy.ascontiguousarray(self._mem)
            if hm.dtype.kind == "u" and hm.dtype.itemsize > 1:
                hm = hm.view("int%d" % (hm.dtype.itemsize * 8))
            src = torch.from_numpy(hm)
            dst = self.devmem_
            dst.copy_(src.reshape(dst.shape).to(dst.device, non_blocking=False))
            self.host_dirty_ = False
            self.dev_dirty_ = False

    def mark_device_written(self):
        """Kernels wrote the device tensor: the host copy is stale until the next map_read."""
        if self.devmem_ is not None:
            self.dev_dirty_ = True
            self.host_dirty_ = False
            log = getattr(SYNC_TRACE, "log", None)
            if log is not None:
                log["marks"].append(self)

    # -- pickling -------------------------------------------------------------------------------------------
    def __getstate__(self):
        self.map_read()
        state = super(Array, self).__getstate__()
        if self.shallow_pickle and self._mem is not None:
            state["_mem"] = None
            state["_shape_dtype"] = (self._mem.shape, str(self._mem.dtype))
        return state

    def __setstate__(self, state):
        sd = state.pop("_shape_dtype", None)
        super(Array, self).__setstate__(state)
        if sd is not None:
            self._mem = numpy.zeros(sd[0], dtype=sd[1])


def _torch_dtype(np_dtype):
    import torch
    return {
        "float32": torch.float32, "float64": torch.float64, "float16": torch.float16,
        "int32": torch.int32, "int64": torch.int64, "uint8": torch.uint8, "int8": torch.int8,
        "int16": torch.int16, "bool": torch.bool, "uint64": torch.int64, "uint32": torch.int32,
        "uint16": torch.int16,
    }[str(numpy.dtype(np_dtype))]


# reference name
Vector = Array
