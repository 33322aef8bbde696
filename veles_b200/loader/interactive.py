"""InteractiveLoader: samples are pushed with ``feed()`` from the user's session or another thread
(reference: ``veles/loader/interactive.py:57-178``).  Used by RESTfulAPI and notebooks; every fed object is
one sample (array, or a path to an image / .npy file)."""
from __future__ import annotations

import queue

import numpy

from ..mutable import Bool
from .base import Loader, TEST, VALID, TRAIN


class InteractiveLoader(Loader):
    """Minibatches are whatever the user feeds from a Python prompt: `feed(array or path)`."""
    MAPPING = "interactive"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(InteractiveLoader, self).__init__(workflow, **kwargs)
        self.sample_shape = tuple(kwargs.get("sample_shape", (784,)))
        self.loadtxt_kwargs = kwargs.get("loadtxt_kwargs", {})
        self.complete = Bool(False)
        self._max_minibatch_size = int(kwargs.get("minibatch_size", 1))
        self.timeout = kwargs.get("timeout", None)

    def init_unpickled(self):
        super(InteractiveLoader, self).init_unpickled()
        self.queue_ = queue.Queue()

    def load_data(self):
        self.class_lengths[TEST] = 1 << 30
        self.class_lengths[VALID] = self.class_lengths[TRAIN] = 0
        self._has_labels = False

    def create_minibatch_data(self):
        self.minibatch_data.reset(numpy.zeros((self._max_minibatch_size,) + self.sample_shape, dtype=numpy.float32))

    def fill_minibatch(self):
        pass

    def analyze_dataset(self):
        pass

    def shuffle(self):
        if not self.shuffled_indices:
            self.shuffled_indices.mem = numpy.zeros(1, dtype=Loader.INDEX_DTYPE)

    def feed(self, obj):
        """Queue one sample (numpy array / nested list / file path); ``None`` ends the session."""
        self.queue_.put(obj)

    def _decode(self, obj):
        if isinstance(obj, str):
            if obj.endswith(".npy"):
                obj = numpy.load(obj)
            elif obj.endswith((".txt", ".csv")):
                obj = numpy.loadtxt(obj, **self.loadtxt_kwargs)
            else:
                from .image import ImagePipeline
                pipe = ImagePipeline(color_space="GRAY" if len(self.sample_shape) < 3 or self.sample_shape[-1] == 1
                                     else "RGB", scale=self.sample_shape[:2] if len(self.sample_shape) >= 2 else 1.0)
                obj = next(pipe.process(pipe.decode(obj)))
        arr = numpy.asarray(obj, dtype=numpy.float32)
        return arr.reshape(self.sample_shape)

    def run(self):
        count = 0
        self.minibatch_data.map_invalidate()
        while count < self._max_minibatch_size:
            try:
                obj = self.queue_.get(timeout=self.timeout) if count == 0 else self.queue_.get_nowait()
            except queue.Empty:
                break
            if obj is None:
                self.complete <<= True
                self.last_minibatch <<= True
                break
            self.minibatch_data.mem[count] = self._decode(obj)
            count += 1
        self.minibatch_data.mem[count:] = 0
        self._minibatch_size_ = count
        self.minibatch_class = TEST
        self.minibatch_data.unmap()
        self.samples_served += count

    def _wrap_run(self, fn):
        from ..units import Unit
        return Unit._wrap_run(self, type(self).run.__get__(self))
