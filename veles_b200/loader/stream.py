"""StreamLoader: minibatches are *fed* from host memory instead of being read from a resident dataset.

It is the B200 counterpart of the reference's host-fed loaders (``InteractiveLoader.feed``,
``veles/loader/interactive.py:57-178``; ``RestfulLoader``, ``veles/loader/restful.py:52-130``) and
of ``Avatar`` double-buffering (``veles/avatar.py:81-95``): ``feed()`` enqueues an asynchronous
copy from (pinned) host memory on a side stream into one of two staging buffers; ``run()`` makes
the compute stream wait for that copy and launches the fused cast+normalise kernel, so the
upload of minibatch i+1 overlaps the training step on minibatch i.
"""
from __future__ import annotations

import numpy

from ..memory import Array
from .base import Loader, TRAIN, VALID, TEST


class StreamLoader(Loader):
    """Minibatches fed by the host program: two pinned staging slots, uploads on their own stream, gather /
    normalise on the device (`Trainer.feed`)."""
    MAPPING = "stream"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(StreamLoader, self).__init__(workflow, **kwargs)
        self.sample_shape = tuple(kwargs.get("sample_shape", (784,)))
        self.data_dtype = kwargs.get("dtype", "float32")
        self.n_classes = int(kwargs.get("n_classes", 10))
        self.samples_per_epoch = int(kwargs.get("samples_per_epoch", 1 << 30))
        self.mean = kwargs.get("mean", None)
        self.rdisp = kwargs.get("rdisp", None)
        # regression targets: None (classification), "input" (autoencoder: the normalised sample is the target)
        # or a shape tuple (targets fed next to the data)
        self.targets = kwargs.get("targets", None)
        if self.targets is not None and self.targets != "input":
            self.targets = tuple(self.targets)
        self._minibatch_targets = Array(shallow_pickle=True)
        self.feed_class = TRAIN

    @property
    def minibatch_targets(self):
        return self._minibatch_targets

    def init_unpickled(self):
        super(StreamLoader, self).init_unpickled()
        self.staging_ = []
        self.staging_labels_ = []
        self.copy_stream_ = None
        self.events_ = []
        self.slot_ = 0
        self.pending_ = []   # [(slot, count)]
        self.identity_ = None
        self.host_pending_ = []

    def load_data(self):
        self.class_lengths[TEST] = 0
        self.class_lengths[VALID] = 0
        self.class_lengths[TRAIN] = self.samples_per_epoch
        self._has_labels = True
        self._labels_mapping = {i: i for i in range(self.n_classes)}
        self._reversed_labels_mapping = list(range(self.n_classes))
        self._unique_labels_count = self.n_classes

    def create_minibatch_data(self):
        self.minibatch_data.reset(numpy.zeros((self._max_minibatch_size,) + self.sample_shape, dtype=numpy.float32))
        if self.targets is not None:
            tshape = (int(numpy.prod(self.sample_shape)),) if self.targets == "input" else self.targets
            self._minibatch_targets.reset(numpy.zeros((self._max_minibatch_size,) + tshape, dtype=numpy.float32))

    def fill_minibatch(self):
        pass

    def analyze_dataset(self):
        pass

    def shuffle(self):
        if not self.shuffled_indices:
            self.shuffled_indices.mem = numpy.zeros(1, dtype=Loader.INDEX_DTYPE)

    def b200_init(self):
        import torch
        dev = self.device_
        self.minibatch_data.dev_dtype = dev.compute_dtype
        self.init_vectors(self.minibatch_data, self.minibatch_labels, self.minibatch_indices)
        self.staging_targets_ = []
        if self.targets is not None:
            self._minibatch_targets.dev_dtype = dev.compute_dtype
            self.init_vectors(self._minibatch_targets)
            if self.targets != "input":
                tsz = int(numpy.prod(self.targets))
                self.staging_targets_ = [torch.zeros(self._max_minibatch_size, tsz, dtype=torch.float32,
                                                     device=dev.torch_device) for _ in range(2)]
        tdt = torch.uint8 if self.data_dtype == "uint8" else torch.float32
        n = self._max_minibatch_size
        size = int(numpy.prod(self.sample_shape))
        self.staging_ = [torch.zeros(n, size, dtype=tdt, device=dev.torch_device) for _ in range(2)]
        self.staging_labels_ = [torch.full((n,), -1, dtype=torch.int32, device=dev.torch_device) for _ in range(2)]
        self.copy_stream_ = torch.cuda.Stream(device=dev.torch_device)
        self.events_ = [torch.cuda.Event() for _ in range(2)]
        self.consumed_ = [torch.cuda.Event() for _ in range(2)]
        for e in self.consumed_:
            e.record()
        self.identity_ = torch.arange(n, dtype=torch.int32, device=dev.torch_device)
        self.mean_dev_ = None if self.mean is None else torch.as_tensor(
            numpy.asarray(self.mean, dtype=numpy.float32).reshape(-1), device=dev.torch_device)
        self.rdisp_dev_ = None if self.rdisp is None else torch.as_tensor(
            numpy.asarray(self.rdisp, dtype=numpy.float32).reshape(-1), device=dev.torch_device)

    # -- feeding -------------------------------------------------------------------------------------------------
    def feed(self, data, labels=None, targets=None):
        """Enqueue one minibatch given as torch CPU tensors (ideally pinned) or numpy arrays.
        Returns the number of bytes that will cross PCIe."""
        import torch
        if isinstance(data, numpy.ndarray):
            data = torch.from_numpy(data)
        if targets is not None and isinstance(targets, numpy.ndarray):
            targets = torch.from_numpy(targets.astype(numpy.float32, copy=False))
        if labels is not None and isinstance(labels, numpy.ndarray):
            labels = torch.from_numpy(labels.astype(numpy.int32, copy=False))
        count = data.shape[0]
        if count > self._max_minibatch_size:
            raise ValueError("feed(): %d samples exceed max_minibatch_size %d" % (count, self._max_minibatch_size))
        if not self.on_device:
            self.host_pending_.append((data, labels, count, targets))
            return 0
        if len(self.pending_) >= 2:
            raise RuntimeError("feed(): both staging slots hold minibatches that were not consumed yet; run a step "
                               "before feeding a third one")
        slot = self.slot_
        self.slot_ ^= 1
        nbytes = data.numel() * data.element_size()
        with torch.cuda.stream(self.copy_stream_):
            # do not overwrite a staging buffer the compute stream may still be reading
            self.copy_stream_.wait_event(self.consumed_[slot])
            self.staging_[slot][:count].copy_(data.reshape(count, -1), non_blocking=True)
            if labels is not None:
                self.staging_labels_[slot][:count].copy_(labels, non_blocking=True)
                nbytes += labels.numel() * labels.element_size()
            if targets is not None and self.staging_targets_:
                self.staging_targets_[slot][:count].copy_(targets.reshape(count, -1), non_blocking=True)
                nbytes += targets.numel() * targets.element_size()
            self.events_[slot].record(self.copy_stream_)
        self.pending_.append((slot, count))
        return nbytes

    def run(self):
        if not self.on_device:
            return self._run_host()
        import torch
        from ..ops import kernels
        if not self.pending_:
            raise RuntimeError("StreamLoader.run(): nothing was fed")
        slot, count = self.pending_.pop(0)
        torch.cuda.current_stream().wait_event(self.events_[slot])
        self._minibatch_size_ = count
        self.minibatch_class = self.feed_class
        md = self.minibatch_data.devmem
        sink = self.__dict__.get("minibatch_sink_") if self.targets is None else None
        if sink is not None and kernels.gather_s2d(self.staging_[slot], self.identity_, 0, count, sink["xs"], sink["H"],
                                                   sink["W"], sink["C"], sink["sy"], sink["sx"], self.mean_dev_,
                                                   self.rdisp_dev_, self.staging_labels_[slot],
                                                   self.minibatch_labels.devmem, self.minibatch_indices.devmem):
            sink["fed"] = True     # the first convolution finds its rearranged input ready
        else:
            if sink is not None:
                sink["fed"] = False
            kernels.gather_minibatch(self.staging_[slot], self.identity_, 0, count, md.view(md.shape[0], -1),
                                     self.mean_dev_, self.rdisp_dev_, self.staging_labels_[slot],
                                     self.minibatch_labels.devmem, self.minibatch_indices.devmem)
        if self.targets is not None:
            mt = self._minibatch_targets.devmem
            mt2 = mt.view(mt.shape[0], -1)
            if self.targets == "input":
                kernels.copy_2d(md.view(md.shape[0], -1), mt2)
            else:
                kernels.gather_minibatch(self.staging_targets_[slot], self.identity_, 0, count, mt2)
            self._minibatch_targets.mark_device_written()
        self.consumed_[slot].record()
        self.samples_served += count
        for a in (self.minibatch_data, self.minibatch_labels, self.minibatch_indices):
            a.mark_device_written()

    def _run_host(self):
        data, labels, count, targets = self.host_pending_.pop(0)
        self._minibatch_size_ = count
        self.minibatch_class = self.feed_class
        arr = data.numpy() if hasattr(data, "numpy") else numpy.asarray(data)
        x = arr.reshape((count,) + self.sample_shape).astype(numpy.float32)
        if self.mean is not None:
            x = (x - numpy.asarray(self.mean, dtype=numpy.float32).reshape(self.sample_shape)) * \
                numpy.asarray(self.rdisp, dtype=numpy.float32).reshape(self.sample_shape)
        self.minibatch_data.mem[:count] = x
        self.minibatch_data.mem[count:] = 0
        self.minibatch_labels.mem[:] = -1
        if labels is not None:
            self.minibatch_labels.mem[:count] = labels.numpy() if hasattr(labels, "numpy") else labels
        if self.targets is not None:
            t = self.minibatch_data.mem.reshape(self.minibatch_data.shape[0], -1) if self.targets == "input" else None
            if t is None:
                t = numpy.zeros(self._minibatch_targets.shape, dtype=numpy.float32)
                if targets is not None:
                    tv = targets.numpy() if hasattr(targets, "numpy") else numpy.asarray(targets)
                    t[:count] = tv.reshape((count,) + tuple(self._minibatch_targets.shape[1:]))
            self._minibatch_targets.mem[...] = t.reshape(self._minibatch_targets.shape)
        self.samples_served += count

    def _wrap_run(self, fn):
        from ..units import Unit
        return Unit._wrap_run(self, type(self).run.__get__(self))
