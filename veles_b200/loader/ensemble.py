"""EnsembleLoader: the features are the stacked outputs of the ensemble members
(reference: ``veles/loader/ensemble.py:53-159``).  Input = the ``--ensemble-test`` result JSON whose models carry
``Output`` [n_samples, n_out] and the dataset labels."""
from __future__ import annotations

import json

import numpy

from .base import TEST, VALID, TRAIN, LoaderError
from .fullbatch import FullBatchLoader, FullBatchLoaderMSEMixin


class EnsembleLoader(FullBatchLoader):
    """Serves the stacked outputs of ensemble members (read from their result files) as features for a
    second-level model."""
    MAPPING = "ensemble"
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(EnsembleLoader, self).__init__(workflow, **kwargs)
        self.file = kwargs["file"]
        self.train_fraction = float(kwargs.get("train_fraction", 0.8))

    def load_data(self):
        if self.original_data and self.restored_from_snapshot:
            return
        with open(self.file) as f:
            data = json.load(f)
        models = [m for m in data["models"] if m and "Output" in m]
        if not models:
            raise LoaderError("%s has no model outputs" % self.file)
        outs = [numpy.asarray(m["Output"], dtype=numpy.float32) for m in models]
        n = outs[0].shape[0]
        if any(o.shape[0] != n for o in outs):
            raise LoaderError("Ensemble members evaluated different numbers of samples")
        self.original_data.mem = numpy.stack(outs, axis=1)     # [n, members, n_out]
        labels = data.get("labels", models[0].get("labels"))
        self.original_labels = list(labels) if labels is not None else []
        n_train = int(n * self.train_fraction)
        self.class_lengths[TEST] = 0
        self.class_lengths[VALID] = n - n_train
        self.class_lengths[TRAIN] = n_train


class EnsembleLoaderMSE(FullBatchLoaderMSEMixin, EnsembleLoader):
    """EnsembleLoader with regression targets."""
    MAPPING = "ensemble_mse"
    hide_from_registry = False

    def load_data(self):
        super(EnsembleLoaderMSE, self).load_data()
        with open(self.file) as f:
            data = json.load(f)
        self.original_targets.mem = numpy.asarray(data["targets"], dtype=numpy.float32)
