"""MinibatchesSaver / MinibatchesLoader: dump served minibatches to one compressed file and replay them.

Parity target: ``veles/loader/saver.py:69-300``: a stream of pickled minibatch chunks per class followed
by a trailing offset table, so the loader can seek to any minibatch; codecs none/gz/bz2/xz."""
from __future__ import annotations

import bz2
import gzip
import lzma
import os
import pickle
import struct

import numpy

from ..config import root, get
from ..distributable import TriviallyDistributable
from ..units import Unit
from .base import Loader, LoaderError, TEST, VALID, TRAIN

CODECS = {"": (lambda b: b, lambda b: b), "gz": (gzip.compress, gzip.decompress),
          "bz2": (bz2.compress, bz2.decompress), "xz": (lzma.compress, lzma.decompress)}
MAGIC = b"VB2MB1"


class MinibatchesSaver(Unit, TriviallyDistributable):
    """Writes every served minibatch to a chunked compressed stream with an offset table (read back by
    MinibatchesLoader)."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "SERVICE")
        super(MinibatchesSaver, self).__init__(workflow, **kwargs)
        self.file_name = os.path.abspath(kwargs.get("file_name", os.path.join(
            get(root.common.dirs.cache, "."), "minibatches.dat")))
        self.compression = kwargs.get("compression", "gz")
        self.class_chunk_sizes = kwargs.get("class_chunk_sizes", (0, 0, 1))
        self.offset_table = []
        self.demand("minibatch_data", "minibatch_class", "class_lengths", "max_minibatch_size", "minibatch_size")
        self.minibatch_labels = None
        self.shuffle_limit = None

    def init_unpickled(self):
        super(MinibatchesSaver, self).init_unpickled()
        self.file_ = None

    def initialize(self, **kwargs):
        os.makedirs(os.path.dirname(self.file_name) or ".", exist_ok=True)
        self.file_ = open(self.file_name, "wb")
        self.file_.write(MAGIC)
        self.offset_table = []

    def run(self):
        n = int(self.minibatch_size)
        self.minibatch_data.map_read()
        labels = None
        if self.minibatch_labels is not None and self.minibatch_labels:
            self.minibatch_labels.map_read()
            labels = self.minibatch_labels.mem[:n].copy()
        payload = CODECS[self.compression][0](pickle.dumps(
            (int(self.minibatch_class), self.minibatch_data.mem[:n].copy(), labels), protocol=4))
        self.offset_table.append((self.file_.tell(), len(payload), int(self.minibatch_class), n))
        self.file_.write(payload)

    def stop(self):
        if self.file_ is None or self.file_.closed:
            return
        table = pickle.dumps({"table": self.offset_table, "compression": self.compression,
                              "class_lengths": list(self.class_lengths),
                              "max_minibatch_size": int(self.max_minibatch_size)}, protocol=4)
        pos = self.file_.tell()
        self.file_.write(table)
        self.file_.write(struct.pack("<QQ", pos, len(table)))
        self.file_.close()
        self.info("Wrote %d minibatches to %s", len(self.offset_table), self.file_name)


class MinibatchesLoader(Loader):
    """Replays a MinibatchesSaver file (samples keep the order in which they were saved)."""
    MAPPING = "minibatches_loader"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(MinibatchesLoader, self).__init__(workflow, **kwargs)
        self.file_name = kwargs["file_name"]
        self.shuffle_limit = 0

    def init_unpickled(self):
        super(MinibatchesLoader, self).init_unpickled()
        self.file_ = None
        self.meta_ = None
        self.index_ = None
        self.cache_ = {}

    def load_data(self):
        self.file_ = open(self.file_name, "rb")
        if self.file_.read(len(MAGIC)) != MAGIC:
            raise LoaderError("%s is not a minibatches file" % self.file_name)
        self.file_.seek(-16, os.SEEK_END)
        pos, size = struct.unpack("<QQ", self.file_.read(16))
        self.file_.seek(pos)
        self.meta_ = pickle.loads(self.file_.read(size))
        # sample index -> (minibatch number, row); samples are laid out class by class
        per_class = [[], [], []]
        for mb, (_, _, klass, n) in enumerate(self.meta_["table"]):
            per_class[klass].extend((mb, r) for r in range(n))
        self.index_ = per_class[TEST] + per_class[VALID] + per_class[TRAIN]
        for ci in range(3):
            self.class_lengths[ci] = len(per_class[ci])
        first = self._read_mb(0)
        self._sample_shape_ = first[1].shape[1:]
        self._has_labels = first[2] is not None
        if self._has_labels:
            labels = set()
            for mb in range(len(self.meta_["table"])):
                labels |= set(int(v) for v in self._read_mb(mb)[2])
            self._labels_mapping = {l: l for l in sorted(labels)}
            self._reversed_labels_mapping = sorted(labels)
            self._unique_labels_count = max(labels) + 1 if labels else 1

    def _read_mb(self, mb):
        if mb not in self.cache_:
            if len(self.cache_) > 64:
                self.cache_.clear()
            off, size, _, _ = self.meta_["table"][mb]
            self.file_.seek(off)
            self.cache_[mb] = pickle.loads(CODECS[self.meta_["compression"]][1](self.file_.read(size)))
        return self.cache_[mb]

    def create_minibatch_data(self):
        self.minibatch_data.reset(numpy.zeros((self._max_minibatch_size,) + tuple(self._sample_shape_),
                                              dtype=numpy.float32))

    def fill_minibatch(self):
        for i, idx in enumerate(self.minibatch_indices.mem[:self._minibatch_size_]):
            mb, row = self.index_[int(idx)]
            _, data, labels = self._read_mb(mb)
            self.minibatch_data.mem[i] = data[row]
            if labels is not None:
                self._raw_minibatch_labels[i] = int(labels[row])
