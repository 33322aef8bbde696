"""Dataset normalizers selected by name.

Parity target: ``veles/normalization.py:110-662``: a registry keyed by ``MAPPING`` with the
names ``none``, ``linear``, ``range_linear``, ``mean_disp``, ``exp``, ``pointwise``,
``external_mean``, ``internal_mean``; two-phase use — ``analyze(data)`` over (chunks of) the
train set, then in-place ``normalize(data)``; ``denormalize``; picklable state.

On the B200 path the ``mean_disp`` coefficients are handed to the gather kernel so the
normalisation is fused into minibatch assembly; everything else runs once on
the host at load time.
"""
from __future__ import annotations

import numpy

from .logger import Logger
from .mapped_object_registry import MappedObjectsRegistry


class UninitializedStateError(Exception):
    pass


class NormalizerRegistry(MappedObjectsRegistry):
    mapping = "normalizers"
    normalizers = {}


def _samples_2d(data):
    """View data as [n_samples, features]."""
    return data.reshape(data.shape[0], -1)


class NormalizerBase(Logger, metaclass=NormalizerRegistry):
    MAPPING = None

    def __init__(self, state=None, **kwargs):
        super(NormalizerBase, self).__init__()
        self._initialized = False
        if state is not None:
            self.state = state

    # subclass API ---------------------------------------------------------------------------------
    def _initialize(self, data):
        pass

    def _analyze(self, data):
        pass

    def _normalize(self, data):
        raise NotImplementedError

    def _denormalize(self, data, **kwargs):
        raise NotImplementedError

    STATE_ATTRS = ()
    STATELESS = False

    # public API -----------------------------------------------------------------------------------
    @property
    def is_initialized(self):
        return self._initialized or self.STATELESS

    def analyze(self, data):
        data = numpy.asarray(data)
        if not self._initialized:
            self._initialize(data)
            self._initialized = True
        self._analyze(data)

    def normalize(self, data):
        if not self.is_initialized:
            raise UninitializedStateError("%s: call analyze() before normalize()" % type(self).__name__)
        return self._normalize(data)

    def denormalize(self, data, **kwargs):
        if not self.is_initialized:
            raise UninitializedStateError("%s: call analyze() before denormalize()" % type(self).__name__)
        return self._denormalize(data, **kwargs)

    def analyze_and_normalize(self, data):
        self.analyze(data)
        return self.normalize(data)

    def reset(self):
        self._initialized = False

    @property
    def state(self):
        if not self.is_initialized:
            raise UninitializedStateError()
        return {k: getattr(self, k) for k in self.STATE_ATTRS}

    @state.setter
    def state(self, value):
        for k, v in value.items():
            setattr(self, k, v)
        self._initialized = True

    @property
    def coefficients(self):
        """(mul, add) such that normalized = data * mul + add, or None when not affine per feature."""
        return None

    def __getstate__(self):
        d = {k: v for k, v in self.__dict__.items() if not k.endswith("_")}
        return d

    def __setstate__(self, state):
        self.__dict__.update(state)


class NoneNormalizer(NormalizerBase):
    MAPPING = "none"
    STATELESS = True

    def _normalize(self, data):
        return None

    def _denormalize(self, data, **kwargs):
        return data


class MeanDispersionNormalizer(NormalizerBase):
    """Subtract the mean and divide by (max - min) per feature (not the statistical dispersion)."""
    MAPPING = "mean_disp"
    STATE_ATTRS = ("_sum", "_count", "_min", "_max")

    def _initialize(self, data):
        self._sum = numpy.zeros_like(data[0], dtype=numpy.float64)
        self._count = 0
        self._min = numpy.array(data[0])
        self._max = numpy.array(data[0])

    def _analyze(self, data):
        self._count += data.shape[0]
        self._sum += numpy.sum(data, axis=0, dtype=numpy.float64)
        numpy.minimum(self._min, data.min(axis=0), out=self._min)
        numpy.maximum(self._max, data.max(axis=0), out=self._max)

    def _coeffs(self):
        mean = self._sum / max(self._count, 1)
        disp = (self._max - self._min).astype(numpy.float64)
        disp[disp == 0] = 1
        return mean, disp

    @property
    def coefficients(self):
        mean, disp = self._coeffs()
        return 1.0 / disp, -mean / disp

    @property
    def mean_rdisp(self):
        mean, disp = self._coeffs()
        return mean.astype(numpy.float32), (1.0 / disp).astype(numpy.float32)

    def _normalize(self, data):
        mean, disp = self._coeffs()
        data -= mean.astype(data.dtype)
        data /= disp.astype(data.dtype)

    def _denormalize(self, data, **kwargs):
        mean, disp = self._coeffs()
        data *= disp.astype(data.dtype)
        data += mean.astype(data.dtype)
        return data


class IntervalMixin(object):
    def _set_interval(self, kwargs):
        self.interval = tuple(kwargs.get("interval", (-1, 1)))
        if len(self.interval) != 2 or self.interval[0] >= self.interval[1]:
            raise ValueError("interval must be (min, max) with min < max")


class LinearNormalizer(NormalizerBase, IntervalMixin):
    """Scale every SAMPLE linearly into ``interval`` using its own min/max (stateless)."""
    MAPPING = "linear"
    STATELESS = True

    def __init__(self, state=None, **kwargs):
        super(LinearNormalizer, self).__init__(state, **kwargs)
        self._set_interval(kwargs)

    def _normalize(self, data):
        flat = _samples_2d(data)
        dmin = flat.min(axis=1, keepdims=True).astype(numpy.float64)
        dmax = flat.max(axis=1, keepdims=True).astype(numpy.float64)
        imin, imax = self.interval
        diff = dmax - dmin
        uniform = diff[:, 0] == 0
        if uniform.any():
            self.warning("There are uniform samples: they are set to the middle of the interval")
        diff[uniform] = 1
        res = (flat - dmin) * ((imax - imin) / diff) + imin
        res[uniform] = (imin + imax) / 2.0
        flat[...] = res.astype(data.dtype)
        return {"dmin": dmin[:, 0], "dmax": dmax[:, 0]}

    def _denormalize(self, data, **kwargs):
        flat = _samples_2d(data)
        dmin = numpy.asarray(kwargs["dmin"], dtype=numpy.float64).reshape(-1, 1)
        dmax = numpy.asarray(kwargs["dmax"], dtype=numpy.float64).reshape(-1, 1)
        imin, imax = self.interval
        flat[...] = ((flat - imin) * ((dmax - dmin) / (imax - imin)) + dmin).astype(data.dtype)
        return data


class RangeLinearNormalizer(NormalizerBase, IntervalMixin):
    """Scale the whole dataset by its GLOBAL min/max into ``interval``."""
    MAPPING = "range_linear"
    STATE_ATTRS = ("_gmin", "_gmax")

    def __init__(self, state=None, **kwargs):
        super(RangeLinearNormalizer, self).__init__(state, **kwargs)
        self._set_interval(kwargs)

    def _initialize(self, data):
        self._gmin = float(data.min())
        self._gmax = float(data.max())

    def _analyze(self, data):
        self._gmin = min(self._gmin, float(data.min()))
        self._gmax = max(self._gmax, float(data.max()))

    @property
    def coefficients(self):
        imin, imax = self.interval
        diff = (self._gmax - self._gmin) or 1.0
        mul = (imax - imin) / diff
        return mul, imin - self._gmin * mul

    def _normalize(self, data):
        mul, add = self.coefficients
        data *= numpy.asarray(mul, dtype=data.dtype)
        data += numpy.asarray(add, dtype=data.dtype)

    def _denormalize(self, data, **kwargs):
        mul, add = self.coefficients
        data -= numpy.asarray(add, dtype=data.dtype)
        data /= numpy.asarray(mul, dtype=data.dtype)
        return data


class ExponentNormalizer(NormalizerBase):
    """Per-sample softmax: exp(x - max) / sum."""
    MAPPING = "exp"
    STATELESS = True

    def _normalize(self, data):
        flat = _samples_2d(data)
        dmax = flat.max(axis=1, keepdims=True)
        flat -= dmax
        numpy.exp(flat, out=flat)
        dsum = flat.sum(axis=1, keepdims=True)
        flat /= dsum
        return {"dmax": dmax[:, 0], "dsum": dsum[:, 0]}

    def _denormalize(self, data, **kwargs):
        flat = _samples_2d(data)
        flat *= numpy.asarray(kwargs["dsum"]).reshape(-1, 1)
        numpy.log(flat, out=flat)
        flat += numpy.asarray(kwargs["dmax"]).reshape(-1, 1)
        return data


class PointwiseNormalizer(NormalizerBase):
    """Per-feature affine map of [min, max] over the train set onto [-1, 1]."""
    MAPPING = "pointwise"
    STATE_ATTRS = ("_min", "_max")

    def _initialize(self, data):
        self._min = data[0].astype(numpy.float64).copy()
        self._max = data[0].astype(numpy.float64).copy()

    def _analyze(self, data):
        numpy.minimum(self._min, data.min(axis=0), out=self._min)
        numpy.maximum(self._max, data.max(axis=0), out=self._max)

    @property
    def coefficients(self):
        disp = self._max - self._min
        mul = numpy.zeros_like(disp)
        nz = disp != 0
        mul[nz] = 2.0 / disp[nz]
        add = numpy.zeros_like(disp)
        add[nz] = -1.0 - self._min[nz] * mul[nz]
        return mul, add

    def _normalize(self, data):
        mul, add = self.coefficients
        data *= mul.astype(data.dtype)
        data += add.astype(data.dtype)

    def _denormalize(self, data, **kwargs):
        mul, add = self.coefficients
        data -= add.astype(data.dtype)
        safe = numpy.where(mul == 0, 1, mul)
        data /= safe.astype(data.dtype)
        return data


class MeanScaleMixin(object):
    def _set_scale(self, kwargs):
        self.scale = float(kwargs.get("scale", 1))


class ExternalMeanNormalizer(NormalizerBase, MeanScaleMixin):
    """Subtract a user-supplied mean sample (array or path to .npy), then multiply by ``scale``."""
    MAPPING = "external_mean"
    STATELESS = True

    def __init__(self, state=None, **kwargs):
        super(ExternalMeanNormalizer, self).__init__(state, **kwargs)
        self._set_scale(kwargs)
        src = kwargs.get("mean_source")
        if src is None and state is None:
            raise ValueError("external_mean normalizer needs mean_source=<array or .npy path>")
        if isinstance(src, str):
            src = numpy.load(src)
        if src is not None:
            self.mean = numpy.asarray(src)

    @property
    def coefficients(self):
        return self.scale, -self.mean * self.scale

    def _normalize(self, data):
        data -= self.mean.astype(data.dtype)
        data *= numpy.asarray(self.scale, dtype=data.dtype)

    def _denormalize(self, data, **kwargs):
        data /= numpy.asarray(self.scale, dtype=data.dtype)
        data += self.mean.astype(data.dtype)
        return data


class InternalMeanNormalizer(NormalizerBase, MeanScaleMixin):
    """Subtract the mean sample computed over the analysed data, then multiply by ``scale``."""
    MAPPING = "internal_mean"
    STATE_ATTRS = ("_sum", "_count")

    def __init__(self, state=None, **kwargs):
        super(InternalMeanNormalizer, self).__init__(state, **kwargs)
        self._set_scale(kwargs)

    def _initialize(self, data):
        self._sum = numpy.zeros_like(data[0], dtype=numpy.float64)
        self._count = 0

    def _analyze(self, data):
        self._sum += data.sum(axis=0, dtype=numpy.float64)
        self._count += data.shape[0]

    @property
    def mean(self):
        return self._sum / max(self._count, 1)

    @property
    def coefficients(self):
        return self.scale, -self.mean * self.scale

    def _normalize(self, data):
        data -= self.mean.astype(data.dtype)
        data *= numpy.asarray(self.scale, dtype=data.dtype)

    def _denormalize(self, data, **kwargs):
        data /= numpy.asarray(self.scale, dtype=data.dtype)
        data += self.mean.astype(data.dtype)
        return data


def factory(name, **kwargs):
    try:
        return NormalizerRegistry.normalizers[name](**kwargs)
    except KeyError:
        raise ValueError("Unknown normalizer \"%s\"; known: %s" % (name, sorted(NormalizerRegistry.normalizers)))
