"""Keyed, picklable, reproducible host random generators.

Parity target: ``veles/prng/random_generator.py:64-303``.  The reference monkey-patches
``numpy.random`` and swaps one global Mersenne state in and out around every call
(``preserve_state :132-140``) so that interleaved users stay reproducible; here every
``RandomGenerator`` simply owns a private ``numpy.random.RandomState`` — same guarantee, no
global patching.  ``seed()`` accepts the CLI forms (bytes / ints / ndarray), ``fill``,
``fill_normal_real`` (clipped at ±5σ, ``:185-198``), ``shuffle``, ``randint``, ... mirror the
reference API.  ``xorshift128plus`` is the CPU reference of the device generator (``:273-283``).
"""
from __future__ import annotations

import threading

import numpy

from ..distributable import Pickleable


class RandomGenerator(Pickleable):
    _lock = threading.Lock()

    def __init__(self, key, **kwargs):
        self._key = key
        self._saved_state = None
        self._seed_value = None
        super(RandomGenerator, self).__init__(**kwargs)

    def init_unpickled(self):
        super(RandomGenerator, self).init_unpickled()
        self.state_ = numpy.random.RandomState()
        if self._saved_state is not None:
            self.state_.set_state(self._saved_state)
        self.lock_ = threading.Lock()

    def __getstate__(self):
        self._saved_state = self.state_.get_state()
        return super(RandomGenerator, self).__getstate__()

    @property
    def key(self):
        return self._key

    @property
    def seed_value(self):
        return self._seed_value

    @property
    def state(self):
        return self.state_.get_state()

    @state.setter
    def state(self, value):
        self.state_.set_state(value)

    def seed(self, seed, dtype=None, count=None):
        """Seed from an int, bytes, an array, or a file path (``count`` items of ``dtype``)."""
        if isinstance(seed, str):
            with open(seed, "rb") as f:
                item = numpy.dtype(dtype or numpy.uint32).itemsize
                seed = numpy.frombuffer(f.read((count or 16) * item), dtype=dtype or numpy.uint32)
        if isinstance(seed, (bytes, bytearray)):
            seed = numpy.frombuffer(bytes(seed) + b"\0" * (-len(seed) % 4), dtype=numpy.uint32)
        if isinstance(seed, numpy.ndarray):
            seed = seed.astype(numpy.uint64) % (2 ** 32)
            seed = seed.astype(numpy.uint32)
        elif seed is not None:
            seed = int(seed) % (2 ** 32)
        self._seed_value = seed
        with self.lock_:
            self.state_.seed(seed)

    # -- draws -----------------------------------------------------------------------------------------
    def _call(self, name, *args, **kwargs):
        with self.lock_:
            return getattr(self.state_, name)(*args, **kwargs)

    def normal(self, loc=0.0, scale=1.0, size=None):
        return self._call("normal", loc, scale, size)

    def rand(self, *args):
        return self._call("rand", *args)

    def random_sample(self, size=None):
        return self._call("random_sample", size)

    def randint(self, low, high=None, size=None):
        return self._call("randint", low, high, size)

    def uniform(self, low=0.0, high=1.0, size=None):
        return self._call("uniform", low, high, size)

    def choice(self, a, size=None, replace=True, p=None):
        return self._call("choice", a, size, replace, p)

    def permutation(self, x):
        return self._call("permutation", x)

    def shuffle(self, arr):
        self._call("shuffle", arr)

    def bytes(self, n):
        return self._call("bytes", n)

    def fill(self, arr, vle_min=-1.0, vle_max=1.0):
        """In-place uniform fill of ``arr`` within [vle_min, vle_max)."""
        flat = arr.reshape(-1)
        flat[:] = (self._call("random_sample", flat.shape) * (vle_max - vle_min) + vle_min).astype(arr.dtype)

    def fill_normal_real(self, arr, mean, stddev, clip_to_sigma=5.0):
        """In-place normal fill clipped to ``clip_to_sigma`` standard deviations."""
        flat = arr.reshape(-1)
        vals = self._call("normal", mean, stddev, flat.shape)
        numpy.clip(vals, mean - clip_to_sigma * stddev, mean + clip_to_sigma * stddev, out=vals)
        flat[:] = vals.astype(arr.dtype)


_generators = {}
_generators_lock = threading.Lock()


def get(key=1):
    """Process-wide generator for ``key`` (created on first use)."""
    with _generators_lock:
        rg = _generators.get(key)
        if rg is None:
            rg = _generators[key] = RandomGenerator(key)
        return rg


def xorshift128plus(states, index):
    """One xorshift128+ step on ``states[index:index+2]`` (uint64), returns the draw."""
    with numpy.errstate(over="ignore"):
        s1 = numpy.uint64(states[index])
        s0 = numpy.uint64(states[index + 1])
        states[index] = s0
        s1 ^= s1 << numpy.uint64(23)
        new = s1 ^ s0 ^ (s1 >> numpy.uint64(17)) ^ (s0 >> numpy.uint64(26))
        states[index + 1] = new
        return numpy.uint64(new + s0)


def xorshift1024star_fill(states, n_rounds):
    """Vectorised numpy oracle of the device xorshift1024* (reference oracle:
    veles/prng/uniform.py:129-163): ``states`` uint64 [S,16] is advanced in place; returns
    uint64 [n_rounds*16*S] with the reference's state-interleaved ordering."""
    S = states.shape[0]
    out = numpy.zeros(n_rounds * 16 * S, dtype=numpy.uint64)
    mult = numpy.uint64(1181783497276652981)
    with numpy.errstate(over="ignore"):
        for r in range(n_rounds):
            for i in range(16):
                s0 = states[:, i].copy()
                s1 = states[:, (i + 1) & 15].copy()
                s1 ^= s1 << numpy.uint64(31)
                s1 ^= s1 >> numpy.uint64(11)
                s0 ^= s0 >> numpy.uint64(30)
                v = s0 ^ s1
                states[:, (i + 1) & 15] = v
                out[(r * 16 + i) * S:(r * 16 + i + 1) * S] = v * mult
    return out
