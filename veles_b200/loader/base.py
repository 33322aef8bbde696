"""Loader base: serves minibatches of a 3-class dataset (test / validation / train).

Parity target: ``veles/loader/base.py:120-1181``.  Kept contract:
  * class order TEST=0, VALID=1, TRAIN=2 (:72-80); ``class_lengths``, ``class_end_offsets``;
  * outputs ``minibatch_data/labels/indices/size/class/offset``, flags ``last_minibatch``,
    ``epoch_ended``, ``train_ended``, ``test_ended`` (Bool), ``epoch_number``, ``samples_served``;
  * subclass contract ``ILoader{load_data, create_minibatch_data, fill_minibatch}`` (:100-154);
  * ``shuffled_indices`` — only the TRAIN part is reshuffled, once per epoch, ``shuffle_limit``
    times (:711-724); ``serve_next_minibatch`` retries failed minibatches first (:726-753);
  * ``--train-ratio`` (:560-566); label mapping and label statistics (:925-1018);
  * job-level master/slave protocol: the master sends (indices, offset, class, size, epoch,
    flags), the slave patches its state (:631-663); ``drop_slave`` re-queues (:679-687).
New: ``shard_rank`` / ``shard_world`` — in synchronous data-parallel runs every rank sees the
same global minibatch sequence and materialises only its contiguous slice, so 1-GPU and
N-GPU runs consume identical sample orders.
"""
from __future__ import annotations

import argparse
import collections
import time

import numpy

from .. import prng
from ..accelerated_units import AcceleratedUnit
from ..config import root, get
from ..error import VelesException
from ..memory import Array
from ..mutable import Bool
from ..normalization import NormalizerRegistry
from ..unit_registry import MappedUnitRegistry
from ..accelerated_units import AcceleratedUnitRegistry
from ..verified import Interface

TARGET = 3
TRAIN = 2
VALID = 1
TEST = 0
CLASS_NAME = ["test", "validation", "train"]


class LoaderError(VelesException):
    pass


class UserLoaderRegistry(AcceleratedUnitRegistry, MappedUnitRegistry):
    mapping = "loaders"
    loaders = {}

    @staticmethod
    def get_factory(name, **kwargs):
        try:
            klass = UserLoaderRegistry.loaders[name]
        except KeyError:
            raise ValueError("Unknown loader \"%s\"; registered: %s"
                             % (name, sorted(UserLoaderRegistry.loaders)))
        return lambda workflow, **kw: klass(workflow, **dict(kwargs, **kw))


class ILoader(Interface):
    def load_data(self):
        """Set class_lengths (and whatever the subclass needs to serve samples)."""

    def create_minibatch_data(self):
        """Allocate minibatch_data for max_minibatch_size samples."""

    def fill_minibatch(self):
        """Fill minibatch_data[:minibatch_size] (and raw labels) from minibatch_indices."""


class Loader(AcceleratedUnit, metaclass=UserLoaderRegistry):
    """Serves data from some source in minibatches."""
    hide_from_registry = True
    LABEL_DTYPE = numpy.int32
    INDEX_DTYPE = numpy.int32
    VIEW_GROUP = "LOADER"

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "LOADER")
        self.last_minibatch = Bool()
        self.epoch_ended = Bool()
        self.train_ended = Bool()
        self.test_ended = Bool()
        super(Loader, self).__init__(workflow, **kwargs)
        self.verify_interface(ILoader)
        self.prng = kwargs.get("prng", prng.get())
        self.shuffle_limit = kwargs.get("shuffle_limit", numpy.iinfo(numpy.uint32).max)
        self._max_minibatch_size = int(kwargs.get("minibatch_size", kwargs.get("max_minibatch_size", 100)))
        if self._max_minibatch_size < 1:
            raise ValueError("minibatch_size must be greater than zero")
        self.class_lengths = [0, 0, 0]
        self.class_end_offsets = [0, 0, 0]
        self._effective_class_end_offsets = [0, 0, 0]
        self._has_labels = False
        self.epoch_number = 0
        self.train_ratio = float(kwargs.get("train_ratio", Loader.cli_train_ratio))
        if not 0.0 < self.train_ratio <= 1.0:
            raise ValueError("train_ratio must be in (0, 1]")
        self.global_offset = 0
        self.minibatch_class = 0
        self.minibatch_data = Array(shallow_pickle=True)
        self.minibatch_indices = Array(shallow_pickle=True)
        self.minibatch_labels = Array(shallow_pickle=True)
        self._raw_minibatch_labels = []
        self._labels_mapping = {}
        self._reversed_labels_mapping = []
        self.failed_minibatches = []
        self._total_failed = 0
        self._unique_labels_count = 1
        self.shuffled_indices = Array()
        self.normalization_type = kwargs.get("normalization_type", "none")
        self.normalization_parameters = kwargs.get("normalization_parameters", {})
        self.testing_mode = kwargs.get("testing", None)
        self._minibatch_offset_ = 0
        self._minibatch_size_ = 0
        self.samples_served = 0
        self.shard_rank = int(kwargs.get("shard_rank", 0))
        self.shard_world = int(kwargs.get("shard_world", 1))
        self.train_only = kwargs.get("train_only", False)

    cli_train_ratio = 1.0

    @staticmethod
    def init_parser(**kwargs):
        parser = kwargs.get("parser", argparse.ArgumentParser())
        parser.add_argument("--train-ratio", default=1.0, type=float,
                            help="Use only this fraction of the train set (ensembles).")
        return parser

    def init_unpickled(self):
        super(Loader, self).init_unpickled()
        self._minibatch_offset_ = 0
        self._minibatch_size_ = 0
        self.pending_minibatches_ = collections.defaultdict(list)
        self._minibatch_serve_timestamp_ = time.time()
        self._normalizer_ = None

    def __getstate__(self):
        state = super(Loader, self).__getstate__()
        # minibatches in flight on slaves are folded back so a resumed master re-serves them
        state["failed_minibatches"] = list(self.failed_minibatches) + \
            [mb for lst in self.pending_minibatches_.values() for mb in lst]
        state["_saved_normalizer"] = self._normalizer_
        return state

    def __setstate__(self, state):
        norm = state.pop("_saved_normalizer", None)
        super(Loader, self).__setstate__(state)
        self._normalizer_ = norm

    # -- properties --------------------------------------------------------------------------------------
    @property
    def has_labels(self):
        return self._has_labels

    @property
    def unique_labels_count(self):
        return self._unique_labels_count

    @property
    def labels_mapping(self):
        return self._labels_mapping

    @property
    def reversed_labels_mapping(self):
        return self._reversed_labels_mapping

    @property
    def max_minibatch_size(self):
        return self._max_minibatch_size

    @max_minibatch_size.setter
    def max_minibatch_size(self, value):
        if value < 1:
            raise ValueError("max_minibatch_size must be greater than zero")
        self._max_minibatch_size = min(int(value), max(self.class_lengths) or int(value))

    @property
    def local_minibatch_size(self):
        """Rows this rank materialises per global minibatch (ceil split)."""
        return -(-self._max_minibatch_size // self.shard_world)

    @property
    def total_samples(self):
        return sum(self.class_lengths)

    @property
    def effective_total_samples(self):
        return self._effective_class_end_offsets[TRAIN]

    @property
    def effective_class_end_offsets(self):
        return self._effective_class_end_offsets

    @property
    def minibatch_offset(self):
        return self._minibatch_offset_

    @minibatch_offset.setter
    def minibatch_offset(self, value):
        self._minibatch_offset_ = int(value)
        self._update_flags()

    @property
    def minibatch_size(self):
        return self._minibatch_size_

    @minibatch_size.setter
    def minibatch_size(self, value):
        if not 0 <= value <= self._max_minibatch_size:
            raise ValueError("Invalid minibatch_size %s (max %d)" % (value, self._max_minibatch_size))
        self._minibatch_size_ = int(value)

    @property
    def normalizer(self):
        if self._normalizer_ is None:
            self._normalizer_ = NormalizerRegistry.normalizers[self.normalization_type](
                **(self.normalization_parameters or {}))
        return self._normalizer_

    @property
    def class_ended(self):
        for offset in self._effective_class_end_offsets:
            if self.global_offset == offset:
                return True
            if self.global_offset < offset:
                return False
        raise LoaderError("global_offset %d is out of bounds %s" % (self.global_offset, self.class_end_offsets))

    @property
    def pending_minibatches_count(self):
        return sum(len(v) for v in self.pending_minibatches_.values())

    @property
    def total_failed(self):
        return self._total_failed

    @property
    def testing_effective(self):
        return self.testing if self.testing_mode is None else bool(self.testing_mode)

    def get_metric_names(self):
        return {"Total epochs"} | ({"Labels"} if self.has_labels else set())

    def get_metric_values(self):
        vals = {"Total epochs": self.epoch_number}
        if self.has_labels:
            vals["Labels"] = self.reversed_labels_mapping
        return vals

    # -- life cycle -----------------------------------------------------------------------------------------
    def initialize(self, device=None, **kwargs):
        self.device = device
        if self.testing_effective:
            self.shuffle_limit = 0
            self.global_offset = 0
        restored = self.restored_from_snapshot and self.total_samples > 0 and bool(self.shuffled_indices)
        try:
            self.load_data()
        except AttributeError:
            self.exception("Failed to load the data")
            raise
        if self.testing_effective and self.class_lengths[TEST] == 0:
            # evaluation mode serves whatever exists as the test set
            self.class_lengths = [sum(self.class_lengths), 0, 0]
        self._calc_class_end_offsets()
        self.info("Samples: test %d, validation %d, train %d (minibatch %d%s)", *(
            tuple(self.class_lengths) + (self._max_minibatch_size,
                                        ", rank %d/%d" % (self.shard_rank, self.shard_world)
                                        if self.shard_world > 1 else "")))
        self._max_minibatch_size = min(self._max_minibatch_size, max(self.class_lengths))
        self.on_before_create_minibatch_data()
        self.create_minibatch_data()
        if not self.minibatch_data:
            raise LoaderError("minibatch_data MUST be initialized in create_minibatch_data()")
        self.analyze_dataset()
        if not restored:
            self.shuffle()
        self._create_label_arrays()
        self._backend_init_()
        return None

    def _create_label_arrays(self):
        n = self._max_minibatch_size
        if not self.minibatch_labels or len(self.minibatch_labels) != n:
            self.minibatch_labels.reset(numpy.full(n, -1, dtype=Loader.LABEL_DTYPE))
        self._raw_minibatch_labels = [None] * n
        if not self.minibatch_indices or len(self.minibatch_indices) != n:
            self.minibatch_indices.reset(numpy.full(n, -1, dtype=Loader.INDEX_DTYPE))

    def on_before_create_minibatch_data(self):
        self.minibatch_data.reset()
        self.minibatch_labels.reset()
        self.minibatch_indices.reset()

    def numpy_init(self):
        pass

    def b200_init(self):
        self.init_vectors(self.minibatch_data, self.minibatch_labels, self.minibatch_indices)

    def run(self):
        """Serve the next minibatch (standalone and slave modes)."""
        if self.is_master:
            return
        self.serve_next_minibatch(None)
        self._on_successful_serve()

    numpy_run = run
    b200_run = run

    def _wrap_run(self, fn):
        # Loader.run is not device-dispatched: keep the Unit wrapper around the plain method
        return super(Loader, self)._wrap_run(type(self).run.__get__(self))

    # -- dataset bookkeeping ---------------------------------------------------------------------------------
    def _calc_class_end_offsets(self):
        total = 0
        for i, n in enumerate(self.class_lengths):
            if n < 0:
                raise LoaderError("class_lengths must be non-negative")
            total += int(n)
            self.class_end_offsets[i] = total
        if total == 0:
            raise LoaderError("There is no data to serve")
        self._effective_class_end_offsets = list(self.class_end_offsets)
        self._effective_class_end_offsets[TRAIN] -= int((1.0 - self.train_ratio) * self.class_lengths[TRAIN])

    def class_index_by_sample_index(self, index):
        for ci, offset in enumerate(self._effective_class_end_offsets):
            if index < offset:
                return ci, offset - index
        raise LoaderError("Sample index %d is out of bounds %s" % (index, self._effective_class_end_offsets))

    def analyze_dataset(self):
        """Feed the train set to the normalizer (host path; FullBatchLoader overrides)."""
        if self.normalization_type == "none" or self.normalizer.is_initialized:
            return
        if self.class_lengths[TRAIN] == 0:
            raise LoaderError("There are no train samples and the normalizer is not initialized")
        self._iterate_class(TRAIN, lambda: self.normalizer.analyze(self.minibatch_data.mem[:self.minibatch_size]))

    def _iterate_class(self, class_index, fn):
        if not self.shuffled_indices:
            self.shuffled_indices.mem = numpy.arange(self.total_samples, dtype=Loader.INDEX_DTYPE)
        if not self.minibatch_indices:
            self.minibatch_indices.reset(numpy.full(self._max_minibatch_size, -1, dtype=Loader.INDEX_DTYPE))
            self._raw_minibatch_labels = [None] * self._max_minibatch_size
        n = self.class_lengths[class_index]
        base = self.class_end_offsets[class_index] - n
        for start in range(0, n, self._max_minibatch_size):
            self._minibatch_size_ = min(self._max_minibatch_size, n - start)
            self.minibatch_indices.mem[:self._minibatch_size_] = \
                self.shuffled_indices.mem[base + start:base + start + self._minibatch_size_]
            self.fill_minibatch()
            fn()

    def shuffle(self):
        """Reshuffle the TRAIN part of shuffled_indices (at most shuffle_limit times)."""
        if not self.shuffled_indices:
            self.shuffled_indices.mem = numpy.arange(self.total_samples, dtype=Loader.INDEX_DTYPE)
        if self.shuffle_limit <= 0 or self.class_lengths[TRAIN] == 0:
            return
        self.shuffle_limit -= 1
        self.shuffled_indices.map_write()
        self.prng.shuffle(self.shuffled_indices.mem[self.class_end_offsets[VALID]:])

    # -- serving ----------------------------------------------------------------------------------------------------
    def serve_next_minibatch(self, slave_id):
        try:
            minibatch_def = self.failed_minibatches.pop()
        except IndexError:
            minibatch_def = self._advance_global_offset()
        minibatch_offset, minibatch_size = minibatch_def
        self.pending_minibatches_[slave_id].append(minibatch_def)
        if not self.is_master and slave_id is None:
            self.pending_minibatches_[slave_id].clear()
        self._minibatch_size_ = minibatch_size
        self.minibatch_offset = minibatch_offset  # updates the flags
        if self.fill_indices(minibatch_offset - minibatch_size, minibatch_size):
            return
        if self.is_master:
            return
        self.fill_minibatch()
        self.normalize_minibatch()
        self.map_minibatch_labels()
        if minibatch_size < self._max_minibatch_size:
            self.minibatch_data.mem[minibatch_size:] = 0
            if self.has_labels:
                self.minibatch_labels.mem[minibatch_size:] = -1
            self.minibatch_indices.mem[minibatch_size:] = -1

    def fill_indices(self, start_offset, count):
        """Host path: copy the index slice; device loaders override and return True."""
        self.minibatch_indices.map_invalidate()
        self.minibatch_indices.mem[:count] = self.shuffled_indices.mem[start_offset:start_offset + count]
        return False

    def normalize_minibatch(self):
        if self.normalization_type != "none":
            self.normalizer.normalize(self.minibatch_data.mem[:self._minibatch_size_])

    def map_minibatch_labels(self):
        if not self.has_labels:
            return
        self.minibatch_labels.map_invalidate()
        for i, raw in enumerate(self._raw_minibatch_labels[:self._minibatch_size_]):
            try:
                self.minibatch_labels.mem[i] = self._labels_mapping[raw]
            except KeyError:
                raise LoaderError("Label %r of sample %d is not in the mapping %s"
                                  % (raw, self.minibatch_indices.mem[i], sorted(self._labels_mapping, key=str)))

    def _update_flags(self):
        if self.is_slave:
            return  # set explicitly in apply_data_from_master
        last_mb = (self.class_ended and (not self.pending_minibatches_count or not self.is_master)
                   and not self.failed_minibatches)
        self.last_minibatch <<= last_mb
        self.epoch_ended <<= last_mb and (
            self.minibatch_class == VALID or
            (self.minibatch_class == TEST and self.class_lengths[TRAIN] == self.class_lengths[VALID] == 0) or
            (self.minibatch_class == TEST and self.testing_effective) or
            (self.minibatch_class == TRAIN and self.class_lengths[VALID] == 0))

    def _advance_global_offset(self):
        if self.is_slave:
            return self.minibatch_offset, self.minibatch_size
        if self.global_offset >= self.effective_total_samples:
            self.global_offset = 0
            self.shuffle()
            self.epoch_number += 1
        if self.train_only and self.global_offset < self.class_end_offsets[VALID] and self.class_lengths[TRAIN]:
            self.global_offset = self.class_end_offsets[VALID]
        self.minibatch_class, remainder = self.class_index_by_sample_index(self.global_offset)
        minibatch_size = min(remainder, self._max_minibatch_size)
        self.global_offset += minibatch_size
        self.train_ended <<= self.global_offset >= self.effective_total_samples
        self.test_ended <<= self.global_offset >= self.class_end_offsets[TEST]
        return self.global_offset, minibatch_size

    def _on_successful_serve(self):
        self.samples_served += self._minibatch_size_
        if not self.is_slave:
            # progress line every 10 s of serving (reference: loader/base.py:399-414)
            now = time.time()
            if now - self.__dict__.get("_serve_log_time_", 0.0) >= 10.0:
                first = "_serve_log_time_" not in self.__dict__
                self._serve_log_time_ = now
                total = max(1, self.effective_total_samples)
                if not first:
                    self.info("Served %d samples (%d epochs, %.1f%% current); jobs failed: %d; pending: %d",
                              self.samples_served, self.samples_served // total,
                              100.0 * (self.samples_served % total) / total, len(self.failed_minibatches),
                              self.pending_minibatches_count)
        if self.last_minibatch:
            self.debug("Last minibatch of class %s served in epoch %d",
                       CLASS_NAME[self.minibatch_class].upper(), self.epoch_number)
            self.event("epoch", "single", number=self.epoch_number, klass=CLASS_NAME[self.minibatch_class])

    # -- labels ----------------------------------------------------------------------------------------------------------
    def _setup_labels_mapping(self, diff_labels):
        """diff_labels: per-class dict label -> count (only classes that exist)."""
        if not self.has_labels:
            return
        other = set()
        for ci in (TEST, VALID):
            other |= set(diff_labels[ci])
        train = set(diff_labels[TRAIN])
        if self.class_lengths[TRAIN] > 0 and other and not other.issubset(train):
            self.warning("There are labels in the test/validation sets absent from the train set: %s",
                         sorted(other - train, key=str))
        for ci in (TEST, VALID, TRAIN):
            if self.class_lengths[ci] > 0:
                self._print_label_stats(diff_labels[ci], CLASS_NAME[ci])
        labels = sorted(train | other, key=lambda v: (str(type(v)), v))
        if not self._labels_mapping or not self.restored_from_snapshot:
            self._labels_mapping = {lbl: i for i, lbl in enumerate(labels)}
            self._reversed_labels_mapping = list(labels)
        self._unique_labels_count = max(len(self._labels_mapping), 1)

    def _print_label_stats(self, stats, set_name):
        """One line per sample class on how evenly its labels are populated (reference: loader/base.py:947-980): a
        spread above 10 % of the mean is flagged, above 50 % flagged as severe; the per-label table goes to DEBUG."""
        values = [int(v) for v in stats.values()]
        if not values or sum(values) == 0:
            self.info("No %s labels specified", set_name)
            return None
        keys = list(stats.keys())
        mean, stddev = float(numpy.mean(values)), float(numpy.std(values))
        lmin, lmax = int(numpy.min(values)), int(numpy.max(values))
        amin, amax = keys[int(numpy.argmin(values))], keys[int(numpy.argmax(values))]
        verdict = "" if stddev <= mean / 10 else (" - UNBALANCED" if stddev > mean / 2 else " - uneven")
        line = ("%s label cardinalities: min: %d (\"%s\"), max: %d (\"%s\"), avg: %d, sigma: %d (%d%%)%s"
                % (set_name, lmin, amin, lmax, amax, int(mean), int(stddev), int(stddev * 100 // max(mean, 1)), verdict))
        # with a handful of samples per label the spread says nothing: a warning only when the counts are meaningful
        (self.warning if verdict == " - UNBALANCED" and mean >= 10 else self.info)(line)
        import logging
        if logging.getLogger(type(self).__name__).isEnabledFor(logging.DEBUG) or \
                logging.getLogger().isEnabledFor(logging.DEBUG):
            total = sum(values)
            rows = ["%-16s %10d %6.1f%%  %s" % (str(k)[:16], v, 100.0 * v / total, "*" * (v * 25 // max(lmax, 1)))
                    for k, v in zip(keys, values)]
            self.debug("Detailed %s label stats:\n%s", set_name, "\n".join(rows))
        return line

    def _validate_and_fix_normalization(self):
        pass

    # -- job-level master/slave protocol ---------------------------------------------------------------------------------
    def generate_data_for_master(self):
        return True

    def generate_data_for_slave(self, slave):
        from ..workflow import NoMoreJobs
        if self.train_ended and not self.failed_minibatches and self._stop_after_epoch(slave):
            raise NoMoreJobs()
        self.serve_next_minibatch(getattr(slave, "id", slave))
        data = {"indices": self.minibatch_indices.mem[:self.minibatch_size].copy(),
                "minibatch_class": self.minibatch_class, "minibatch_size": self.minibatch_size,
                "minibatch_offset": self.minibatch_offset, "epoch_number": self.epoch_number,
                "last_minibatch": bool(self.last_minibatch), "epoch_ended": bool(self.epoch_ended),
                "train_ended": bool(self.train_ended), "test_ended": bool(self.test_ended)}
        # epoch boundary = sync point: wait for every outstanding update before serving on
        if self.class_ended:
            self.has_data_for_slave = False   # re-enabled when the last outstanding update has arrived
        return data

    def _stop_after_epoch(self, slave):
        return False

    def apply_data_from_master(self, data):
        for attr in ("minibatch_class", "epoch_number"):
            setattr(self, attr, data[attr])
        self._minibatch_size_ = data["minibatch_size"]
        self._minibatch_offset_ = data["minibatch_offset"]
        self.last_minibatch <<= data["last_minibatch"]
        self.epoch_ended <<= data["epoch_ended"]
        self.train_ended <<= data["train_ended"]
        self.test_ended <<= data.get("test_ended", False)
        indices = data["indices"]
        if indices.size != self._minibatch_size_:
            raise LoaderError("Job indices size %d != minibatch_size %d" % (indices.size, self._minibatch_size_))
        # patch the local view of the global shuffle so fill_indices works unchanged
        start = self._minibatch_offset_ - self._minibatch_size_
        if not self.shuffled_indices:
            self.shuffled_indices.mem = numpy.arange(self.total_samples, dtype=Loader.INDEX_DTYPE)
        self.shuffled_indices.map_write()
        self.shuffled_indices.mem[start:start + self._minibatch_size_] = indices

    def apply_data_from_slave(self, data, slave):
        sid = getattr(slave, "id", slave)
        if sid is None:
            return
        try:
            self.pending_minibatches_[sid].pop(0)
        except (KeyError, IndexError):
            self.warning("Update from %s without a pending minibatch", sid)
        if not self.pending_minibatches_.get(sid):
            self.pending_minibatches_.pop(sid, None)
        self._update_flags()
        self._on_successful_serve()
        if not self.pending_minibatches_count:
            self.has_data_for_slave = True

    def drop_slave(self, slave):
        sid = getattr(slave, "id", slave)
        pending = self.pending_minibatches_.pop(sid, [])
        if pending:
            self._total_failed += len(pending)
            self.failed_minibatches.extend(pending)
            self.info("Jobs failed: %d / pending: %d", len(self.failed_minibatches), self.pending_minibatches_count)
        self.has_data_for_slave = True


class LoaderMSEMixin(object):
    """Adds regression targets (reference: veles/loader/base.py:1034-1146)."""
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(LoaderMSEMixin, self).__init__(workflow, **kwargs)
        self.class_targets = Array()
        self._minibatch_targets = Array(shallow_pickle=True)
        self._targets_shape = kwargs.get("targets_shape", tuple())
        self.target_normalization_type = kwargs.get("target_normalization_type",
                                                    kwargs.get("normalization_type", "none"))
        self.target_normalization_parameters = kwargs.get("target_normalization_parameters",
                                                          kwargs.get("normalization_parameters", {}))
        self._target_normalizer_ = None

    @property
    def minibatch_targets(self):
        return self._minibatch_targets

    @property
    def targets_shape(self):
        return self._targets_shape

    @property
    def target_normalizer(self):
        if self._target_normalizer_ is None:
            self._target_normalizer_ = NormalizerRegistry.normalizers[self.target_normalization_type](
                **(self.target_normalization_parameters or {}))
        return self._target_normalizer_

    def on_before_create_minibatch_data(self):
        super(LoaderMSEMixin, self).on_before_create_minibatch_data()
        self._minibatch_targets.reset()

    def serve_next_minibatch(self, slave_id):
        super(LoaderMSEMixin, self).serve_next_minibatch(slave_id)
        if self._minibatch_targets.devmem is None and self._minibatch_size_ < self._max_minibatch_size \
                and self._minibatch_targets:
            self._minibatch_targets.mem[self._minibatch_size_:] = 0


class LoaderMSE(LoaderMSEMixin, Loader):
    hide_from_registry = True


class LoaderWithValidationRatio(Loader):
    """Splits a validation set off the train set (reference :1158-1181)."""
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(LoaderWithValidationRatio, self).__init__(workflow, **kwargs)
        self.validation_ratio = kwargs.get("validation_ratio", None)

    @property
    def validation_ratio(self):
        return self._validation_ratio

    @validation_ratio.setter
    def validation_ratio(self, value):
        if value is not None and not 0 <= value < 1:
            raise ValueError("validation_ratio must be in [0, 1)")
        self._validation_ratio = value
