"""HDF5 full-batch loader (reference: ``veles/loader/loader_hdf5.py``).  ``h5py`` is optional: without it the
loader raises a clear error at ``load_data`` time, and ``.npz`` archives with the same keys are accepted."""
from __future__ import annotations

import numpy

from .base import TEST, VALID, TRAIN, LoaderError
from .fullbatch import FullBatchLoader


class HDF5Loader(FullBatchLoader):
    """Samples and labels from HDF5 files (needs h5py)."""
    MAPPING = "full_batch_hdf5"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(HDF5Loader, self).__init__(workflow, **kwargs)
        self.test_path = kwargs.get("test_path")
        self.validation_path = kwargs.get("validation_path")
        self.train_path = kwargs.get("train_path")
        self.data_key = kwargs.get("data_key", "data")
        self.labels_key = kwargs.get("labels_key", "label")

    def _open(self, path):
        if path.endswith(".npz"):
            arch = numpy.load(path)
            return numpy.asarray(arch[self.data_key]), numpy.asarray(arch[self.labels_key]).reshape(-1)
        try:
            import h5py
        except ImportError:
            raise LoaderError("h5py is not installed: convert %s to .npz (keys %r, %r) or install h5py"
                              % (path, self.data_key, self.labels_key))
        with h5py.File(path, "r") as f:
            return numpy.asarray(f[self.data_key]), numpy.asarray(f[self.labels_key]).reshape(-1)

    def load_data(self):
        if self.original_data and self.restored_from_snapshot:
            return
        datas, labels = [], []
        for ci, path in ((TEST, self.test_path), (VALID, self.validation_path), (TRAIN, self.train_path)):
            if not path:
                continue
            d, l = self._open(path)
            datas.append(d)
            labels.extend(int(v) for v in l)
            self.class_lengths[ci] = len(l)
        if not datas:
            raise LoaderError("No HDF5 files were given")
        self.original_data.mem = numpy.concatenate(datas)
        self.original_labels = labels
