"""Decision units: per-class epoch statistics, improvement tracking, stop criterion.

``DecisionGD`` / ``DecisionMSE`` (``docs/source/manualrst_veles_workflow_parameters.rst:
141-144``; kwargs ``fail_iterations``, ``max_epochs`` ``:613-614``).  They publish
``improved`` (gates the Snapshotter), ``complete`` (gates the loop / EndPoint), ``gd_skip``
(gates the GD units on non-train minibatches), ``snapshot_suffix`` (e.g.
``validation_1.48_train_0.00``, cf. ``docs/source/manualrst_veles_algorithms.rst:31``) and the
``EvaluationFitness`` result consumed by genetics / ensembles
(``veles/genetics/optimization_workflow.py:248``, ``veles/ensemble/model_workflow.py:132``).

B200 design: evaluator counters stay on the device and are fetched only when a class ends
(``last_minibatch``) — one small D2H read per class per epoch instead of one per minibatch;
in data-parallel runs the counters are summed over ranks at that point.
"""
from __future__ import annotations

import time

import numpy

from ..distributable import TriviallyDistributable
from ..loader.base import TEST, VALID, TRAIN, CLASS_NAME
from ..memory import Array
from ..mutable import Bool
from ..units import Unit


class DecisionBase(Unit):
    hide_from_registry = True
    VIEW_GROUP = "TRAINER"

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "TRAINER")
        super(DecisionBase, self).__init__(workflow, **kwargs)
        self.max_epochs = kwargs.get("max_epochs", None)
        self.fail_iterations = kwargs.get("fail_iterations", 100)
        self.improved = Bool(False)
        self.train_improved = Bool(False)
        self.complete = Bool(False)
        self.gd_skip = Bool(False)
        self.epoch_ended = None  # linked from the loader (shared Bool)
        self.snapshot_suffix = ""
        self.epoch_number = None
        self.minibatch_class = None
        self.minibatch_size = None
        self.last_minibatch = None
        self.class_lengths = None
        self.epoch_timestamp = None
        self._epochs_without_improvement = 0
        self.epoch_metrics_history = []
        self.demand("minibatch_class", "last_minibatch", "class_lengths", "epoch_number", "epoch_ended")
        self.dp_group = None  # set by the data-parallel launcher: sums counters over ranks

    def init_unpickled(self):
        super(DecisionBase, self).init_unpickled()
        self._epoch_started_ = time.time()

    def initialize(self, **kwargs):
        if self.epoch_timestamp is None:
            self.epoch_timestamp = time.time()
        if self.testing:
            # evaluation of a snapshot: one pass over the test set, whatever state training stopped in (a snapshot
            # written in the last epoch carries complete = True, which would keep the loader's gate shut)
            self.max_epochs = None
            self.complete <<= False
            self.improved <<= False
        elif bool(self.complete) and self.class_lengths is not None and self.epoch_number is not None:
            # resumed from the snapshot of a finished run: go on if this run's limits allow more epochs
            done = int(self.epoch_number) + (0 if self.has_validation else 1)
            if ((self.max_epochs is None or done < self.max_epochs)
                    and self._epochs_without_improvement < self.fail_iterations):
                self.complete <<= False
        return None

    @property
    def has_validation(self):
        return self.class_lengths[VALID] > 0

    @property
    def decisive_class(self):
        if self.has_validation:
            return VALID
        return TRAIN if self.class_lengths[TRAIN] > 0 else TEST

    def run(self):
        klass = int(self.minibatch_class)
        self.gd_skip <<= (klass != TRAIN) or self.testing
        if self.is_slave:
            # a slave job is ONE minibatch: close the loop, the master accumulates the statistics
            self.complete <<= True
            return
        self.on_minibatch(klass)
        if not self.last_minibatch:
            return
        self.on_class_ended(klass)
        # the loader raises epoch_ended on the last VALIDATION minibatch (or the last train /
        # test one when there is no validation set): reference veles/loader/base.py:862-878
        if self.epoch_ended:
            self.on_epoch_ended()

    def on_minibatch(self, klass):
        pass

    def on_class_ended(self, klass):
        pass

    def on_epoch_ended(self):
        stop = False
        if self.testing:
            stop = True
        done_epochs = int(self.epoch_number) + (0 if self.has_validation else 1)
        if self.max_epochs is not None and done_epochs >= self.max_epochs:
            stop = True
        if self.class_lengths[TRAIN] > 0 and self._epochs_without_improvement >= self.fail_iterations:
            stop = True
        if self.class_lengths[TRAIN] == 0:
            stop = True
        self.event("epoch", "end", number=int(self.epoch_number))
        self.epoch_timestamp = time.time()
        if stop:
            self.complete <<= True

    # -- job-level master/slave: the slave reports per-minibatch statistics, the master owns epochs ---------
    def generate_data_for_slave(self, slave):
        return {"epoch_number": int(self.epoch_number)}

    def apply_data_from_master(self, data):
        self.complete <<= False
        self.improved <<= False

    def generate_data_for_master(self):
        data = {"minibatch_class": int(self.minibatch_class), "last_minibatch": bool(self.last_minibatch),
                "epoch_ended": bool(self.epoch_ended), "stats": self.collect_minibatch_stats()}
        return data

    def apply_data_from_slave(self, data, slave):
        klass = data["minibatch_class"]
        self.merge_minibatch_stats(klass, data["stats"])
        # the MASTER's loader flags (already refreshed by Loader.apply_data_from_slave, which precedes this
        # unit) say whether that was the last outstanding minibatch of the class / epoch
        if self.last_minibatch:
            self.on_class_ended(klass)
            if self.epoch_ended:
                self.on_epoch_ended()

    def collect_minibatch_stats(self):
        return None

    def merge_minibatch_stats(self, klass, stats):
        pass

    def drop_slave(self, slave):
        pass


class DecisionGD(DecisionBase):
    """Classification: tracks the number of errors per class."""
    hide_from_registry = False
    MAPPING = "decision_gd"
    LOSS = "softmax"

    def __init__(self, workflow, **kwargs):
        super(DecisionGD, self).__init__(workflow, **kwargs)
        self.epoch_n_err = [0, 0, 0]
        self.epoch_n_evaluated = [0, 0, 0]
        self.epoch_n_err_pt = [100.0, 100.0, 100.0]
        self.best_n_err_pt = [100.0, 100.0, 100.0]
        self.best_epoch = 0
        self.epoch_loss = [0.0, 0.0, 0.0]
        self.minibatch_n_err = None          # evaluator.n_err ([wrong, evaluated], accumulating)
        self.minibatch_confusion_matrix = None
        self.minibatch_max_err_y_sum = None
        self.minibatch_loss_sum = None
        self.confusion_matrixes = [None, None, None]
        self.max_err_y_sums = [0.0, 0.0, 0.0]
        self.demand("minibatch_n_err")

    def _counters_on_device(self):
        return isinstance(self.minibatch_n_err, Array) and self.minibatch_n_err.devmem is not None

    def on_minibatch(self, klass):
        # host counters are cheap to drain every minibatch; device counters wait for the class end
        if not self._counters_on_device():
            self._drain(klass)

    def _drain(self, klass):
        ne = self.minibatch_n_err
        ne.map_read()
        vals = numpy.array(ne.mem[:2], dtype=numpy.int64)
        loss = 0.0
        if self.minibatch_loss_sum is not None and self.minibatch_loss_sum:
            self.minibatch_loss_sum.map_read()
            loss = float(self.minibatch_loss_sum.mem[0])
        if self.dp_group is not None:
            vals, loss = self.dp_group.sum_counters(vals, loss)
        self.epoch_n_err[klass] += int(vals[0])
        self.epoch_n_evaluated[klass] += int(vals[1])
        self.epoch_loss[klass] += loss
        ne.map_write()
        ne.mem[:2] = 0
        ne.unmap()
        if self.minibatch_loss_sum is not None and self.minibatch_loss_sum:
            self.minibatch_loss_sum.map_write()
            self.minibatch_loss_sum.mem[0] = 0
            self.minibatch_loss_sum.unmap()
        cm = self.minibatch_confusion_matrix
        if cm is not None and cm:
            cm.map_read()
            if self.confusion_matrixes[klass] is None:
                self.confusion_matrixes[klass] = numpy.zeros_like(cm.mem)
            # every rank evaluated its shard of the class: the matrix of the whole class is the sum over ranks
            self.confusion_matrixes[klass] += (cm.mem if self.dp_group is None else self.dp_group.sum_array(cm.mem))
            cm.map_write()
            cm.mem[...] = 0
            cm.unmap()
        ms = self.minibatch_max_err_y_sum
        if ms is not None and ms:
            ms.map_read()
            top = float(ms.mem[0]) if self.dp_group is None else float(self.dp_group.sum_array(ms.mem[:1], op="max")[0])
            self.max_err_y_sums[klass] = max(self.max_err_y_sums[klass], top)
            ms.map_write()
            ms.mem[0] = 0
            ms.unmap()

    def collect_minibatch_stats(self):
        saved = (list(self.epoch_n_err), list(self.epoch_n_evaluated), list(self.epoch_loss))
        self.epoch_n_err, self.epoch_n_evaluated, self.epoch_loss = [0, 0, 0], [0, 0, 0], [0.0, 0.0, 0.0]
        klass = int(self.minibatch_class)
        self._drain(klass)
        stats = (self.epoch_n_err[klass], self.epoch_n_evaluated[klass], self.epoch_loss[klass])
        self.epoch_n_err, self.epoch_n_evaluated, self.epoch_loss = saved
        return stats

    def merge_minibatch_stats(self, klass, stats):
        self.epoch_n_err[klass] += stats[0]
        self.epoch_n_evaluated[klass] += stats[1]
        self.epoch_loss[klass] += stats[2]

    def on_class_ended(self, klass):
        if self._counters_on_device() and not self.is_master:
            self._drain(klass)
        n = max(self.epoch_n_evaluated[klass], 1)
        self.epoch_n_err_pt[klass] = 100.0 * self.epoch_n_err[klass] / n
        if klass == self.decisive_class:
            better = self.epoch_n_err_pt[klass] < self.best_n_err_pt[klass]
            self.improved <<= better
            if better:
                self.best_n_err_pt[klass] = self.epoch_n_err_pt[klass]
                self.best_epoch = int(self.epoch_number)
                self._epochs_without_improvement = 0
            else:
                self._epochs_without_improvement += 1
            self.snapshot_suffix = "validation_%.2f_train_%.2f" % (
                self.epoch_n_err_pt[VALID], self.epoch_n_err_pt[TRAIN])
        if klass == TRAIN:
            tb = self.epoch_n_err_pt[TRAIN] < self.best_n_err_pt[TRAIN] or not self.has_validation
            self.train_improved <<= tb
            if self.has_validation and self.epoch_n_err_pt[TRAIN] < self.best_n_err_pt[TRAIN]:
                self.best_n_err_pt[TRAIN] = self.epoch_n_err_pt[TRAIN]

    def on_epoch_ended(self):
        self.info("Epoch %d: errors %s (%s%%), loss %s%s", int(self.epoch_number),
                  self.epoch_n_err, ", ".join("%.2f" % v for v in self.epoch_n_err_pt),
                  ", ".join("%.4f" % (self.epoch_loss[c] / max(self.epoch_n_evaluated[c], 1)) for c in range(3)),
                  " *" if self.improved else "")
        self.epoch_metrics_history.append({"epoch": int(self.epoch_number), "n_err": list(self.epoch_n_err),
                                           "n_err_pt": list(self.epoch_n_err_pt),
                                           "loss": [self.epoch_loss[c] / max(self.epoch_n_evaluated[c], 1)
                                                    for c in range(3)]})
        super(DecisionGD, self).on_epoch_ended()
        self.epoch_n_err = [0, 0, 0]
        self.epoch_n_evaluated = [0, 0, 0]
        self.epoch_loss = [0.0, 0.0, 0.0]

    def get_metric_names(self):
        return {"EvaluationFitness", "Errors", "Best epoch", "Confusion"}

    def get_metric_values(self):
        k = self.decisive_class
        return {"EvaluationFitness": 1.0 - self.best_n_err_pt[k] / 100.0,
                "Errors": {CLASS_NAME[c]: self.best_n_err_pt[c] for c in range(3)},
                "Best epoch": self.best_epoch,
                "Confusion": {CLASS_NAME[c]: (m.tolist() if m is not None else None)
                              for c, m in enumerate(self.confusion_matrixes)}}


class DecisionMSE(DecisionBase):
    """Regression: tracks mean / max / min RMSE per class."""
    hide_from_registry = False
    MAPPING = "decision_mse"
    LOSS = "mse"

    def __init__(self, workflow, **kwargs):
        super(DecisionMSE, self).__init__(workflow, **kwargs)
        self.epoch_metrics = [[0.0, 0.0, 1e30] for _ in range(3)]
        self.epoch_n_evaluated = [0, 0, 0]
        self.epoch_rmse = [1e30, 1e30, 1e30]
        self.best_rmse = [1e30, 1e30, 1e30]
        self.best_epoch = 0
        self.minibatch_metrics = None   # evaluator.metrics [sum, max, min], accumulating
        self.minibatch_n_err = None     # evaluator.n_err ([_, evaluated])
        self.demand("minibatch_metrics", "minibatch_n_err")

    def _on_device(self):
        return isinstance(self.minibatch_metrics, Array) and self.minibatch_metrics.devmem is not None

    def on_minibatch(self, klass):
        if not self._on_device():
            self._drain(klass)

    def _drain(self, klass):
        mm, ne = self.minibatch_metrics, self.minibatch_n_err
        mm.map_read()
        ne.map_read()
        s, mx, mn = float(mm.mem[0]), float(mm.mem[1]), float(mm.mem[2])
        cnt = int(ne.mem[1])
        if self.dp_group is not None:
            vals, s = self.dp_group.sum_counters(numpy.array([0, cnt], dtype=numpy.int64), s)
            cnt = int(vals[1])
            mx, mn = self.dp_group.max_min(mx, mn)
        e = self.epoch_metrics[klass]
        e[0] += s
        e[1] = max(e[1], mx)
        e[2] = min(e[2], mn)
        self.epoch_n_evaluated[klass] += cnt
        mm.map_write()
        mm.mem[...] = [0, 0, 1e30]
        mm.unmap()
        ne.map_write()
        ne.mem[:2] = 0
        ne.unmap()

    def collect_minibatch_stats(self):
        saved = ([list(m) for m in self.epoch_metrics], list(self.epoch_n_evaluated))
        self.epoch_metrics = [[0.0, 0.0, 1e30] for _ in range(3)]
        self.epoch_n_evaluated = [0, 0, 0]
        klass = int(self.minibatch_class)
        self._drain(klass)
        stats = (list(self.epoch_metrics[klass]), self.epoch_n_evaluated[klass])
        self.epoch_metrics, self.epoch_n_evaluated = saved
        return stats

    def merge_minibatch_stats(self, klass, stats):
        e = self.epoch_metrics[klass]
        e[0] += stats[0][0]
        e[1] = max(e[1], stats[0][1])
        e[2] = min(e[2], stats[0][2])
        self.epoch_n_evaluated[klass] += stats[1]

    def on_class_ended(self, klass):
        if self._on_device() and not self.is_master:
            self._drain(klass)
        self.epoch_rmse[klass] = self.epoch_metrics[klass][0] / max(self.epoch_n_evaluated[klass], 1)
        if klass == self.decisive_class:
            better = self.epoch_rmse[klass] < self.best_rmse[klass]
            self.improved <<= better
            if better:
                self.best_rmse[klass] = self.epoch_rmse[klass]
                self.best_epoch = int(self.epoch_number)
                self._epochs_without_improvement = 0
            else:
                self._epochs_without_improvement += 1
            self.snapshot_suffix = "validation_%.4f_train_%.4f" % (
                min(self.epoch_rmse[VALID], 9999), min(self.epoch_rmse[TRAIN], 9999))

    def on_epoch_ended(self):
        self.info("Epoch %d: rmse %s%s", int(self.epoch_number),
                  ", ".join("%.4f" % min(v, 9999) for v in self.epoch_rmse), " *" if self.improved else "")
        self.epoch_metrics_history.append({"epoch": int(self.epoch_number), "rmse": list(self.epoch_rmse)})
        super(DecisionMSE, self).on_epoch_ended()
        self.epoch_metrics = [[0.0, 0.0, 1e30] for _ in range(3)]
        self.epoch_n_evaluated = [0, 0, 0]

    def get_metric_names(self):
        return {"EvaluationFitness", "RMSE", "Best epoch"}

    def get_metric_values(self):
        k = self.decisive_class
        return {"EvaluationFitness": -self.best_rmse[k], "RMSE": {CLASS_NAME[c]: self.best_rmse[c] for c in range(3)},
                "Best epoch": self.best_epoch}
