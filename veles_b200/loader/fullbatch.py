"""FullBatchLoader: the whole dataset lives in one array, device-resident when it fits.

Parity target: ``veles/loader/fullbatch.py:79-566``: ``original_data`` / ``original_labels``
(/ ``original_targets``) hold every sample; minibatches are assembled **on the device** by a
gather kernel (``:197-221, 292-310``; reference kernel ``cuda/fullbatch_loader.cu:9-61``);
whole-dataset normalisation happens once at initialisation (``:336-347``); OOM falls back to
host-resident data (``:168-187``); ``validation_ratio`` carves a stratified validation set
out of train (``:349-433``).

B200 design: one fused kernel (``vb_gather_minibatch``) gathers by shuffled index, casts
(uint8/fp32 -> bf16/fp32), applies the mean/dispersion normalisation when the normalizer is
affine per feature, pads the tail with zeros / -1 labels, and materialises only this rank's
slice of the global minibatch.  With 180 GB of HBM3e the dataset stays resident in its
*storage* dtype (uint8 images are not expanded to float).
"""
from __future__ import annotations

import numpy

from ..memory import Array
from ..verified import Interface
from .base import Loader, LoaderMSEMixin, LoaderError, TEST, VALID, TRAIN


class IFullBatchLoader(Interface):
    def load_data(self):
        """Fill original_data (and original_labels / class_lengths)."""


class FullBatchLoader(Loader):
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(FullBatchLoader, self).__init__(workflow, **kwargs)
        self.validation_ratio = kwargs.get("validation_ratio", None)
        self.original_data = Array()
        self.original_labels = []
        self._mapped_original_labels_ = Array()
        self.force_numpy_data = kwargs.get("force_numpy_data", False)
        self.store_in_device_memory = kwargs.get("store_in_device_memory", True)
        self.fused_normalization = kwargs.get("fused_normalization", True)

    def init_unpickled(self):
        super(FullBatchLoader, self).init_unpickled()
        self._mapped_original_labels_ = Array()
        self._mean_dev_ = None
        self._rdisp_dev_ = None
        self._shuffled_dev_ = None
        self._shuffled_version_ = -1
        self._shuffle_count_ = 0

    @property
    def on_device_data(self):
        return self.on_device and self.original_data.devmem is not None

    # -- ILoader ---------------------------------------------------------------------------------------
    def create_minibatch_data(self):
        rows = self.local_minibatch_size if self.shard_world > 1 else self._max_minibatch_size
        self.minibatch_data.reset(numpy.zeros((rows,) + self.original_data.shape[1:], dtype=numpy.float32))

    def _create_label_arrays(self):
        super(FullBatchLoader, self)._create_label_arrays()
        if self.shard_world > 1:
            n = self.local_minibatch_size
            self.minibatch_labels.reset(numpy.full(n, -1, dtype=Loader.LABEL_DTYPE))
            self.minibatch_indices.reset(numpy.full(n, -1, dtype=Loader.INDEX_DTYPE))

    def fill_minibatch(self):
        idx = self.minibatch_indices.mem[:self._minibatch_size_]
        self.minibatch_data.mem[:self._minibatch_size_] = self.original_data.mem[idx]
        if self.has_labels:
            for i, ind in enumerate(idx):
                self._raw_minibatch_labels[i] = self.original_labels[int(ind)]

    def normalize_minibatch(self):
        pass  # the whole dataset was normalised once in analyze_dataset()

    def analyze_dataset(self):
        """Normalise the whole dataset once (host), or keep affine coefficients for the fused kernel."""
        self._setup_labels()
        if self.normalization_type == "none":
            return
        data = self.original_data.mem
        norm = self.normalizer
        if not norm.is_initialized:
            if self.class_lengths[TRAIN] > 0:
                train = data[self.class_end_offsets[VALID]:]
                step = max(1, 65536 * 64 // max(1, self.original_data.sample_size))
                for i in range(0, len(train), step):
                    norm.analyze(train[i:i + step].astype(numpy.float32, copy=False))
            elif not getattr(norm, "STATELESS", False):
                raise LoaderError("There are no train samples and the normalizer is not initialized")
        fuse = (self.fused_normalization and self.on_device and not self.force_numpy_data
                and norm.coefficients is not None)
        if fuse:
            mul, add = norm.coefficients
            mul = numpy.broadcast_to(numpy.asarray(mul, dtype=numpy.float64), data.shape[1:]).reshape(-1)
            add = numpy.broadcast_to(numpy.asarray(add, dtype=numpy.float64), data.shape[1:]).reshape(-1)
            safe = numpy.where(mul == 0, 1.0, mul)
            # kernel form: (x - mean) * rdisp  with  rdisp = mul, mean = -add / mul
            self._fused_mean_rdisp = ((-add / safe).astype(numpy.float32), mul.astype(numpy.float32))
            return
        self._fused_mean_rdisp = None
        if data.dtype.kind != "f":
            self.original_data.mem = data = data.astype(numpy.float32)
        for i in range(0, len(data), 4096):
            norm.normalize(data[i:i + 4096])

    _fused_mean_rdisp = None

    def _setup_labels(self):
        self._has_labels = len(self.original_labels) > 0
        if not self._has_labels:
            return
        if len(self.original_labels) != self.total_samples:
            raise LoaderError("original_labels has %d entries for %d samples"
                              % (len(self.original_labels), self.total_samples))
        diff = []
        start = 0
        for ci in range(3):
            end = self.class_end_offsets[ci]
            cnt = {}
            for l in self.original_labels[start:end]:
                cnt[l] = cnt.get(l, 0) + 1
            diff.append(cnt)
            start = end
        self._setup_labels_mapping(diff)
        mapped = numpy.fromiter((self._labels_mapping[l] for l in self.original_labels),
                                dtype=Loader.LABEL_DTYPE, count=len(self.original_labels))
        self._mapped_original_labels_.reset(mapped)

    def initialize(self, device=None, **kwargs):
        res = super(FullBatchLoader, self).initialize(device=device, **kwargs)
        return res

    def load_data(self):
        raise NotImplementedError

    # -- validation split ---------------------------------------------------------------------------------
    def resize_validation(self, ratio=None):
        """Move a stratified ``ratio`` of train samples into the validation class."""
        ratio = self.validation_ratio if ratio is None else ratio
        if ratio is None or ratio <= 0:
            return
        n_test, n_valid, n_train = self.class_lengths
        pool_start = n_test
        pool = numpy.arange(pool_start, pool_start + n_valid + n_train)
        labels = self.original_labels
        rng = self.prng
        if labels:
            by_label = {}
            for i in pool:
                by_label.setdefault(labels[int(i)], []).append(int(i))
            valid_idx = []
            for lbl in sorted(by_label, key=str):
                lst = numpy.array(by_label[lbl])
                rng.shuffle(lst)
                valid_idx.extend(lst[:int(round(len(lst) * ratio))].tolist())
        else:
            perm = pool.copy()
            rng.shuffle(perm)
            valid_idx = perm[:int(round(len(pool) * ratio))].tolist()
        valid_set = set(valid_idx)
        train_idx = [int(i) for i in pool if int(i) not in valid_set]
        order = list(range(n_test)) + sorted(valid_idx) + train_idx
        self.original_data.mem = self.original_data.mem[order]
        if labels:
            self.original_labels = [labels[i] for i in order]
        tgt = getattr(self, "original_targets", None)
        if tgt is not None and tgt:
            tgt.mem = tgt.mem[order]
        self.class_lengths[VALID] = len(valid_idx)
        self.class_lengths[TRAIN] = len(train_idx)
        self._calc_class_end_offsets()

    # -- device path ------------------------------------------------------------------------------------------
    def b200_init(self):
        import torch
        dev = self.device_
        self.minibatch_data.dev_dtype = dev.compute_dtype
        self.init_vectors(self.minibatch_data, self.minibatch_labels, self.minibatch_indices)
        if self.force_numpy_data or not self.store_in_device_memory:
            return
        try:
            self.original_data.initialize(dev)
            if self._has_labels:
                self._mapped_original_labels_.initialize(dev)
        except (RuntimeError, torch.cuda.OutOfMemoryError) as e:  # dataset does not fit: host path
            self.warning("Failed to place the dataset on the device (%s): serving from the host", e)
            self.original_data._free_device()
            self.force_numpy_data = True
            return
        if self._fused_mean_rdisp is not None:
            mean, rdisp = self._fused_mean_rdisp
            self._mean_dev_ = torch.from_numpy(mean).to(dev.torch_device)
            self._rdisp_dev_ = torch.from_numpy(rdisp).to(dev.torch_device)

    def shuffle(self):
        super(FullBatchLoader, self).shuffle()
        self._shuffle_count_ = getattr(self, "_shuffle_count_", 0) + 1

    def _device_indices(self):
        import torch
        if self._shuffled_dev_ is None or self._shuffled_version_ != self._shuffle_count_ \
                or self.is_slave:
            self._shuffled_dev_ = torch.from_numpy(
                numpy.ascontiguousarray(self.shuffled_indices.mem.astype(numpy.int32))).to(
                    self.device_.torch_device, non_blocking=True)
            self._shuffled_version_ = self._shuffle_count_
        return self._shuffled_dev_

    def _fill_shard_host(self, start_offset, count):
        """numpy backend, data parallel: materialise this rank's contiguous slice of the global minibatch."""
        per = self.local_minibatch_size
        lo = min(count, self.shard_rank * per)
        hi = min(count, lo + per)
        n = hi - lo
        self.local_count = n
        idx = self.shuffled_indices.mem[start_offset + lo:start_offset + hi]
        for a in (self.minibatch_indices, self.minibatch_data, self.minibatch_labels):
            if a:
                a.map_invalidate()
        self.minibatch_indices.mem[:n] = idx
        self.minibatch_indices.mem[n:] = -1
        self.minibatch_data.mem[:n] = self.original_data.mem[idx]
        self.minibatch_data.mem[n:] = 0
        if self._has_labels:
            self.minibatch_labels.mem[:n] = self._mapped_original_labels_.mem[idx]
            self.minibatch_labels.mem[n:] = -1
        targets = getattr(self, "original_targets", None)
        if targets is not None and targets:
            self._minibatch_targets.map_invalidate()
            self._minibatch_targets.mem[:n] = targets.mem[idx]
            self._minibatch_targets.mem[n:] = 0
        return True

    def fill_indices(self, start_offset, count):
        if not self.on_device_data or self.force_numpy_data:
            if self.shard_world > 1:
                return self._fill_shard_host(start_offset, count)
            return super(FullBatchLoader, self).fill_indices(start_offset, count)
        from ..ops import kernels
        # this rank's contiguous slice of the global minibatch
        if self.shard_world > 1:
            per = self.local_minibatch_size
            lo = min(count, self.shard_rank * per)
            hi = min(count, lo + per)
            start_offset, count = start_offset + lo, hi - lo
        self.local_count = count
        idx = self._device_indices()
        data2d = self.minibatch_data.devmem
        lab = (self._mapped_original_labels_.devmem if self._has_labels else None,
               self.minibatch_labels.devmem if self._has_labels else None, self.minibatch_indices.devmem)
        sink = self.__dict__.get("minibatch_sink_")
        if sink is not None and kernels.gather_s2d(self.original_data.devmem, idx, start_offset, count, sink["xs"],
                                                   sink["H"], sink["W"], sink["C"], sink["sy"], sink["sx"],
                                                   self._mean_dev_, self._rdisp_dev_, *lab):
            sink["fed"] = True     # the consumer finds its rearranged input ready (minibatch_data is not written)
        else:
            if sink is not None:
                sink["fed"] = False
            kernels.gather_minibatch(self.original_data.devmem, idx, start_offset, count, data2d,
                                     self._mean_dev_, self._rdisp_dev_, *lab)
        self.minibatch_data.mark_device_written()
        self.minibatch_labels.mark_device_written()
        self.minibatch_indices.mark_device_written()
        return True

    local_count = 0


class FullBatchLoaderMSEMixin(LoaderMSEMixin):
    """Adds ``original_targets`` gathered by the same kernel (reference :467-566)."""
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        super(FullBatchLoaderMSEMixin, self).__init__(workflow, **kwargs)
        self.original_targets = Array()

    def create_minibatch_data(self):
        super(FullBatchLoaderMSEMixin, self).create_minibatch_data()
        rows = self.minibatch_data.shape[0]
        self._minibatch_targets.reset(numpy.zeros((rows,) + self.original_targets.shape[1:], dtype=numpy.float32))

    def analyze_dataset(self):
        super(FullBatchLoaderMSEMixin, self).analyze_dataset()
        if self.target_normalization_type != "none" and self.original_targets:
            t = self.original_targets.mem
            if t.dtype.kind != "f":
                self.original_targets.mem = t = t.astype(numpy.float32)
            tn = self.target_normalizer
            if not tn.is_initialized:
                tn.analyze(t[self.class_end_offsets[VALID]:] if self.class_lengths[TRAIN] else t)
            tn.normalize(t)

    def fill_minibatch(self):
        super(FullBatchLoaderMSEMixin, self).fill_minibatch()
        idx = self.minibatch_indices.mem[:self._minibatch_size_]
        self._minibatch_targets.mem[:self._minibatch_size_] = self.original_targets.mem[idx]

    def b200_init(self):
        super(FullBatchLoaderMSEMixin, self).b200_init()
        self._minibatch_targets.dev_dtype = self.device_.compute_dtype
        self._minibatch_targets._pitched = True
        self.init_vectors(self._minibatch_targets)
        if not self.force_numpy_data and self.store_in_device_memory:
            self.original_targets.initialize(self.device_)

    def fill_indices(self, start_offset, count):
        done = super(FullBatchLoaderMSEMixin, self).fill_indices(start_offset, count)
        if done and self.on_device_data and not self.force_numpy_data:   # (the host shard path fills targets itself)
            from ..ops import kernels
            if self.shard_world > 1:
                per = self.local_minibatch_size
                lo = min(count, self.shard_rank * per)
                start_offset, count = start_offset + lo, min(count, lo + per) - lo
            kernels.gather_minibatch(self.original_targets.devmem, self._device_indices(), start_offset, count,
                                     self._minibatch_targets.devmem)
            self._minibatch_targets.mark_device_written()
        return done


class FullBatchLoaderMSE(FullBatchLoaderMSEMixin, FullBatchLoader):
    hide_from_registry = True
