"""CIFAR-style pickled batches: ``full_batch_pickles_image`` (reference: ``veles/loader/pickles.py:55-215``).
Each pickle holds a dict with ``data`` [n, C*H*W] (or [n,H,W,C]) and ``labels``."""
from __future__ import annotations

import pickle

import numpy

from .base import TEST, VALID, TRAIN, LoaderError
from .fullbatch import FullBatchLoader


class PicklesImageFullBatchLoader(FullBatchLoader):
    """Whole data set from pickle files of (images, labels), kept resident."""
    MAPPING = "full_batch_pickles_image"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(PicklesImageFullBatchLoader, self).__init__(workflow, **kwargs)
        self.test_pickles = list(kwargs.get("test_pickles", []))
        self.validation_pickles = list(kwargs.get("validation_pickles", []))
        self.train_pickles = list(kwargs.get("train_pickles", []))
        self.image_shape = tuple(kwargs.get("image_shape", (32, 32, 3)))   # HWC
        self.planar = kwargs.get("planar", True)                            # CIFAR stores CHW planes
        self.data_key = kwargs.get("data_key", "data")
        self.labels_key = kwargs.get("labels_key", "labels")

    def _read(self, path):
        with open(path, "rb") as f:
            d = pickle.load(f, encoding="latin1")
        data = numpy.asarray(d[self.data_key] if self.data_key in d else d[self.data_key.encode()])
        labels = d.get(self.labels_key, d.get(self.labels_key.encode(), d.get("fine_labels")))
        n = data.shape[0]
        h, w, c = self.image_shape
        if data.ndim == 2:
            data = data.reshape(n, c, h, w).transpose(0, 2, 3, 1) if self.planar else data.reshape(n, h, w, c)
        return numpy.ascontiguousarray(data), list(labels)

    def load_data(self):
        if self.original_data and self.restored_from_snapshot:
            return
        datas, labels = [], []
        for ci, paths in ((TEST, self.test_pickles), (VALID, self.validation_pickles), (TRAIN, self.train_pickles)):
            count = 0
            for p in paths:
                d, l = self._read(p)
                datas.append(d)
                labels.extend(l)
                count += len(l)
            self.class_lengths[ci] = count
        if not datas:
            raise LoaderError("No pickles were given")
        self.original_data.mem = numpy.concatenate(datas)
        self.original_labels = [int(v) for v in labels]
        if self.validation_ratio:
            self.resize_validation()
