"""Evaluators: turn network output + ground truth into ``err_output`` and metrics.

``EvaluatorSoftmax`` / ``EvaluatorMSE`` (``docs/source/manualrst_veles_workflow_parameters.rst:
163-166``; kwargs ``root``, ``mean``, ``compute_confusion_matrix`` ``:648-650``).  Pinned
semantics: softmax ``err_output = (y - onehot(label)) / batch_size`` for labelled rows and 0
for padded rows (label -1); ``n_err`` = [wrong, evaluated]; ``confusion_matrix[pred, truth]``;
``max_err_output_sum`` = max over samples of Σ|err_output|.  MSE: ``err_output = (y - target)
/ batch_size``; ``metrics`` = [Σ rmse, max rmse, min rmse].

Device path: ONE kernel does softmax + argmax + evaluator from the pre-softmax sums
(csrc/nn_kernels.cu ``softmax_eval_kernel``); counters accumulate on the device and are read
by the Decision unit only at class/epoch boundaries, so the hot loop never synchronises.
"""
from __future__ import annotations

import numpy

from ..accelerated_units import AcceleratedUnit
from ..distributable import TriviallyDistributable
from ..memory import Array


class EvaluatorBase(AcceleratedUnit, TriviallyDistributable):
    hide_from_registry = True
    VIEW_GROUP = "EVALUATOR"

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "EVALUATOR")
        super(EvaluatorBase, self).__init__(workflow, **kwargs)
        self.mean = kwargs.get("mean", True)
        self.err_output = Array(shallow_pickle=True)
        self.output = None
        self.batch_size = None
        self.testing_mode = False
        self.world_size = int(kwargs.get("world_size", 1))
        self.demand("output", "batch_size")

    @property
    def err_scale(self):
        bs = int(self.batch_size) if self.batch_size else 1
        return 1.0 / max(bs, 1) if self.mean else 1.0


class EvaluatorSoftmax(EvaluatorBase):
    """Cross-entropy gradient, error count, loss sum, confusion matrix and max error from softmax outputs and
    labels; counters accumulate on the device."""
    hide_from_registry = False
    MAPPING = "evaluator_softmax"
    LOSS = "softmax"

    def __init__(self, workflow, **kwargs):
        super(EvaluatorSoftmax, self).__init__(workflow, **kwargs)
        self.compute_confusion_matrix = kwargs.get("compute_confusion_matrix", True)
        self.confusion_matrix = Array()
        self.n_err = Array()
        self.max_err_output_sum = Array()
        self.loss_sum = Array()
        self.labels = None
        self.max_idx = None
        self.softmax_unit = kwargs.get("softmax_unit", None)
        self.demand("labels")

    def initialize(self, device=None, **kwargs):
        if self.output is None or not self.output:
            return True
        self.device = device
        n_classes = self.output.size // self.output.shape[0]
        if not self.err_output or self.err_output.shape != self.output.shape:
            self.err_output.reset(numpy.zeros(self.output.shape, dtype=numpy.float32))
        if not self.n_err:
            self.n_err.reset(numpy.zeros(2, dtype=numpy.int32))
        if not self.max_err_output_sum:
            self.max_err_output_sum.reset(numpy.zeros(1, dtype=numpy.float32))
        if not self.loss_sum:
            self.loss_sum.reset(numpy.zeros(1, dtype=numpy.float32))
        if self.compute_confusion_matrix and (not self.confusion_matrix or
                                              self.confusion_matrix.shape != (n_classes, n_classes)):
            self.confusion_matrix.reset(numpy.zeros((n_classes, n_classes), dtype=numpy.int32))
        self._backend_init_()
        return None

    def numpy_run(self):
        self.output.map_read()
        self.labels.map_read()
        y = self.output.mem.reshape(self.output.shape[0], -1).astype(numpy.float32)
        labels = self.labels.mem
        rows = min(len(labels), y.shape[0])
        valid = labels[:rows] >= 0
        idx = numpy.nonzero(valid)[0]
        err = numpy.zeros_like(y)
        err[idx] = y[idx]
        err[idx, labels[idx]] -= 1.0
        err *= self.err_scale
        self.err_output.map_invalidate()
        self.err_output.mem[...] = err.reshape(self.err_output.shape)
        pred = y.argmax(axis=1)
        self.n_err.map_write()
        self.n_err.mem[0] += int((pred[idx] != labels[idx]).sum())
        self.n_err.mem[1] += len(idx)
        if self.compute_confusion_matrix:
            self.confusion_matrix.map_write()
            numpy.add.at(self.confusion_matrix.mem, (pred[idx], labels[idx]), 1)
        self.max_err_output_sum.map_write()
        if len(idx):
            self.max_err_output_sum.mem[0] = max(self.max_err_output_sum.mem[0], numpy.abs(err[idx]).sum(axis=1).max())
            self.loss_sum.map_write()
            self.loss_sum.mem[0] += float(-numpy.log(numpy.maximum(y[idx, labels[idx]], 1e-30)).sum())

    def b200_init(self):
        dev = self.device_
        self.err_output.dev_dtype = dev.compute_dtype
        self.err_output._pitched = True
        self.init_vectors(self.err_output, self.n_err, self.max_err_output_sum, self.loss_sum)
        if self.compute_confusion_matrix:
            self.confusion_matrix.initialize(dev)
        sm = self.softmax_unit
        if sm is not None and getattr(sm, "on_device", False):
            sm.defer_softmax = True

    def b200_run(self):
        from ..ops import kernels
        sm = self.softmax_unit
        if sm is None or getattr(sm, "logits_", None) is None:
            # no pre-softmax sums available: evaluate on the host (off the hot path)
            self.numpy_run()
            for a in (self.err_output, self.n_err, self.max_err_output_sum, self.loss_sum, self.confusion_matrix):
                a.unmap()
            return
        self.labels.unmap()
        out = self.output.devmem
        out2 = out if out.dim() == 2 else out.reshape(out.shape[0], -1)
        err = self.err_output.devmem
        err2 = err if err.dim() == 2 else err.reshape(err.shape[0], -1)
        kernels.softmax_eval(sm.logits_, out2, err2, self.labels.devmem, sm.max_idx.devmem,
                             err_scale=self.err_scale, n_err=self.n_err.devmem,
                             confusion=self.confusion_matrix.devmem if self.compute_confusion_matrix else None,
                             max_err_sum=self.max_err_output_sum.devmem, loss_sum=self.loss_sum.devmem)
        for a in (self.output, self.err_output, self.n_err, self.max_err_output_sum, self.loss_sum, sm.max_idx):
            a.mark_device_written()
        if self.compute_confusion_matrix:
            self.confusion_matrix.mark_device_written()


class EvaluatorMSE(EvaluatorBase):
    """Squared-error gradient and metrics (sum, max, min of per-sample MSE) from outputs and targets."""
    hide_from_registry = False
    MAPPING = "evaluator_mse"
    LOSS = "mse"

    def __init__(self, workflow, **kwargs):
        super(EvaluatorMSE, self).__init__(workflow, **kwargs)
        self.target = None
        self.metrics = Array()
        self.mse = Array(shallow_pickle=True)
        self.n_err = Array()
        self.root = kwargs.get("root", True)
        self.labels = None
        self.class_targets = None
        self.demand("target")

    def initialize(self, device=None, **kwargs):
        if self.output is None or not self.output or self.target is None or not self.target:
            return True
        self.device = device
        if not self.err_output or self.err_output.shape != self.output.shape:
            self.err_output.reset(numpy.zeros(self.output.shape, dtype=numpy.float32))
        if not self.metrics:
            self.metrics.reset(numpy.array([0, 0, 1.0e30], dtype=numpy.float32))
        if not self.mse or len(self.mse) != self.output.shape[0]:
            self.mse.reset(numpy.zeros(self.output.shape[0], dtype=numpy.float32))
        if not self.n_err:
            self.n_err.reset(numpy.zeros(2, dtype=numpy.int32))
        self._backend_init_()
        return None

    def numpy_run(self):
        self.output.map_read()
        self.target.map_read()
        y = self.output.mem.reshape(self.output.shape[0], -1).astype(numpy.float32)
        t = self.target.mem.reshape(self.target.shape[0], -1).astype(numpy.float32)
        valid = min(int(self.batch_size), y.shape[0]) if self.batch_size is not None else y.shape[0]
        local = getattr(self, "local_batch_size", None)
        if local is not None:
            valid = min(int(local), y.shape[0])
        e = numpy.zeros_like(y)
        e[:valid] = y[:valid] - t[:valid]
        rm = numpy.sqrt((e * e).mean(axis=1)) if self.root else (e * e).mean(axis=1)
        self.err_output.map_invalidate()
        self.err_output.mem[...] = (e * self.err_scale).reshape(self.err_output.shape)
        self.mse.map_invalidate()
        self.mse.mem[...] = rm
        self.metrics.map_write()
        if valid:
            self.metrics.mem[0] += rm[:valid].sum()
            self.metrics.mem[1] = max(self.metrics.mem[1], rm[:valid].max())
            self.metrics.mem[2] = min(self.metrics.mem[2], rm[:valid].min())
        self.n_err.map_write()
        self.n_err.mem[1] += valid

    def b200_init(self):
        dev = self.device_
        self.err_output.dev_dtype = dev.compute_dtype
        self.err_output._pitched = True
        self.init_vectors(self.err_output, self.metrics, self.mse, self.n_err)

    def b200_run(self):
        import torch
        from ..ops import kernels
        y = self.output.devmem
        y2 = y if y.dim() == 2 else y.reshape(y.shape[0], -1)
        t = self.target.devmem
        t2 = t if t.dim() == 2 else t.reshape(t.shape[0], -1)
        if t2.dtype != y2.dtype:
            if getattr(self, "_t_stage_", None) is None:
                self._t_stage_ = kernels.pitched_empty(t2.shape[0], t2.shape[1], y2.dtype, y2.device)
            kernels.copy_2d(t2, self._t_stage_)
            t2 = self._t_stage_
        err = self.err_output.devmem
        err2 = err if err.dim() == 2 else err.reshape(err.shape[0], -1)
        local = getattr(self, "local_batch_size", None)
        valid = int(local) if local is not None else min(int(self.batch_size), y2.shape[0])
        kernels.mse_eval(y2, t2, err2, valid, err_scale=self.err_scale, metrics=self.metrics.devmem,
                         mse_per_sample=self.mse.devmem)
        self.n_err.devmem[1] += valid
        for a in (self.err_output, self.metrics, self.mse, self.n_err):
            a.mark_device_written()
