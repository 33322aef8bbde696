"""Numpy statements of the signal primitives: the oracle the native (C++) and device (CUDA) paths are tested against.

Semantics follow the reference's libSimd (``libVeles/simd/src/{convolve,correlate,wavelet,normalize,detect_peaks}.c``);
every function takes one signal (1-D) or a batch of rows (2-D) and works in float32 like the reference does.
"""
from __future__ import annotations

import json
import os

import numpy

WAVELET_DAUBECHIES, WAVELET_COIFLET, WAVELET_SYMLET = 0, 1, 2
EXT_PERIODIC, EXT_MIRROR, EXT_CONSTANT, EXT_ZERO = 0, 1, 2, 3
EXTREMUM_MAXIMUM, EXTREMUM_MINIMUM, EXTREMUM_BOTH = 1, 2, 3
CONV_BRUTE, CONV_FFT, CONV_OVERLAP_SAVE = 0, 1, 2

_TYPE_NAMES = {WAVELET_DAUBECHIES: "daubechies", WAVELET_COIFLET: "coiflet", WAVELET_SYMLET: "symlet",
               "daubechies": "daubechies", "coiflet": "coiflet", "symlet": "symlet",
               "db": "daubechies", "coif": "coiflet", "sym": "symlet"}
_TYPE_IDS = {"daubechies": WAVELET_DAUBECHIES, "coiflet": WAVELET_COIFLET, "symlet": WAVELET_SYMLET}
_TABLES = None


def wavelet_tables():
    """{family: {order: float64 taps}} - orthonormal scaling filters computed by scripts/gen_wavelet_tables.py."""
    global _TABLES
    if _TABLES is None:
        with open(os.path.join(os.path.dirname(__file__), "wavelet_tables.json")) as f:
            raw = json.load(f)
        _TABLES = {fam: {int(o): numpy.asarray(v, dtype=numpy.float64) for o, v in tab.items()}
                   for fam, tab in raw.items()}
    return _TABLES


def wavelet_type_name(wtype):
    try:
        return _TYPE_NAMES[wtype]
    except KeyError:
        raise ValueError("unknown wavelet type %r" % (wtype,))


def wavelet_type_id(wtype):
    return _TYPE_IDS[wavelet_type_name(wtype)]


def wavelet_validate_order(wtype, order):
    return int(order) in wavelet_tables()[wavelet_type_name(wtype)]


def wavelet_filters(wtype, order, reference_scale=True):
    """(lowpass, highpass) float32 taps as the transform applies them.  ``reference_scale`` reproduces the
    reference's tables: Daubechies sums to sqrt 2, coiflets and symlets to 1 (``simd/src/symlets.c:1594``)."""
    fam = wavelet_type_name(wtype)
    if not wavelet_validate_order(fam, order):
        raise ValueError("%s wavelet with %d taps is not available" % (fam, order))
    taps = wavelet_tables()[fam][int(order)]
    if reference_scale and fam != "daubechies":
        taps = taps / numpy.sqrt(2.0)
    lo = taps.astype(numpy.float32)
    hi = numpy.empty_like(lo)
    idx = numpy.arange(order)
    hi[order - 1 - idx] = numpy.where(idx & 1, lo, -lo)
    return lo, hi


def _extend(src, extra, ext):
    n = src.shape[-1]
    i = numpy.arange(extra)
    if ext == EXT_PERIODIC:
        tail = src[..., i % n]
    elif ext == EXT_MIRROR:
        tail = src[..., n - 1 - (i % n)]
    elif ext == EXT_CONSTANT:
        tail = numpy.repeat(src[..., n - 1:n], extra, axis=-1)
    elif ext == EXT_ZERO:
        tail = numpy.zeros(src.shape[:-1] + (extra,), dtype=src.dtype)
    else:
        raise ValueError("unknown extension %r" % (ext,))
    return numpy.concatenate([src, tail], axis=-1)


def wavelet_apply(wtype, order, ext, src, reference_scale=True):
    """One decimated analysis step: (hi, lo), each ``length / 2`` long (``wavelet.c:271-322``)."""
    src = numpy.asarray(src, dtype=numpy.float32)
    n = src.shape[-1]
    if n % 2:
        raise ValueError("the decimated transform needs an even length")
    lo_t, hi_t = wavelet_filters(wtype, order, reference_scale)
    xe = _extend(src, order, ext)
    win = numpy.lib.stride_tricks.sliding_window_view(xe, order, axis=-1)[..., 0:n:2, :]
    return ((win * hi_t).sum(-1, dtype=numpy.float32).astype(numpy.float32),
            (win * lo_t).sum(-1, dtype=numpy.float32).astype(numpy.float32))


def stationary_wavelet_apply(wtype, order, level, ext, src, reference_scale=True):
    """One undecimated step at ``level`` (taps ``2^(level-1)`` apart): (hi, lo) of the input length
    (``wavelet.c:325-392``)."""
    src = numpy.asarray(src, dtype=numpy.float32)
    n = src.shape[-1]
    stride = 1 << (level - 1)
    lo_t, hi_t = wavelet_filters(wtype, order, reference_scale)
    xe = _extend(src, order * stride, ext)
    hi = numpy.zeros(src.shape, dtype=numpy.float32)
    lo = numpy.zeros(src.shape, dtype=numpy.float32)
    for j in range(order):
        seg = xe[..., j * stride:j * stride + n]
        hi += hi_t[j] * seg
        lo += lo_t[j] * seg
    return hi, lo


def convolve(x, h):
    """Full linear convolution, ``len(x) + len(h) - 1`` outputs (``convolve.c:40-101``)."""
    x = numpy.asarray(x, dtype=numpy.float32)
    h = numpy.asarray(h, dtype=numpy.float32)
    if x.ndim == 1:
        return numpy.convolve(x.astype(numpy.float64), h.astype(numpy.float64)).astype(numpy.float32)
    hh = numpy.broadcast_to(h, (x.shape[0], h.shape[-1])) if h.ndim == 1 else h
    return numpy.stack([convolve(a, b) for a, b in zip(x, hh)])


def cross_correlate(x, h):
    """``result[k] = sum_m x[m] h[m + len(h) - 1 - k]`` = ``numpy.correlate(x, h, "full")`` (``correlate.c:74-130``)."""
    x = numpy.asarray(x, dtype=numpy.float32)
    h = numpy.asarray(h, dtype=numpy.float32)
    if x.ndim == 1:
        return numpy.correlate(x.astype(numpy.float64), h.astype(numpy.float64), "full").astype(numpy.float32)
    hh = numpy.broadcast_to(h, (x.shape[0], h.shape[-1])) if h.ndim == 1 else h
    return numpy.stack([cross_correlate(a, b) for a, b in zip(x, hh)])


def convolve_choose(x_length, h_length):
    """The reference's algorithm heuristic (``convolve.c:347-363``)."""
    if x_length > 2 * h_length:
        return CONV_OVERLAP_SAVE if x_length > 200 else CONV_BRUTE
    return CONV_FFT if x_length > 350 else CONV_BRUTE


def minmax1D(src):
    src = numpy.asarray(src, dtype=numpy.float32)
    return src.min(axis=-1), src.max(axis=-1)


def minmax2D(src):
    src = numpy.asarray(src, dtype=numpy.uint8)
    return src.min(axis=(-2, -1)), src.max(axis=(-2, -1))


def normalize2D_minmax(vmin, vmax, src):
    """``(v - min) / ((max - min) / 2) - 1`` per plane; a flat plane gives zeros (``normalize.c:382-400``)."""
    src = numpy.asarray(src, dtype=numpy.uint8)
    vmin = numpy.asarray(vmin, dtype=numpy.float32)[..., None, None]
    vmax = numpy.asarray(vmax, dtype=numpy.float32)[..., None, None]
    diff = (vmax - vmin) / 2.0
    safe = numpy.where(diff > 0, diff, 1.0)
    out = (src.astype(numpy.float32) - vmin) / safe - 1.0
    return numpy.where(diff > 0, out, 0.0).astype(numpy.float32)


def normalize2D(src):
    lo, hi = minmax2D(src)
    return normalize2D_minmax(lo, hi, src)


def detect_peaks(data, extremum_type=EXTREMUM_BOTH):
    """Strict interior extrema in index order: (positions int32, values float32) (``detect_peaks.c:43-60``)."""
    data = numpy.asarray(data, dtype=numpy.float32)
    if data.ndim != 1:
        raise ValueError("the oracle takes one signal; loop over a batch")
    if data.size < 3:
        return numpy.zeros(0, numpy.int32), numpy.zeros(0, numpy.float32)
    d1 = data[1:-1] - data[:-2]
    d2 = data[1:-1] - data[2:]
    flag = (d1 * d2 > 0) & (((d1 > 0) & bool(extremum_type & EXTREMUM_MAXIMUM)) |
                            ((d1 < 0) & bool(extremum_type & EXTREMUM_MINIMUM)))
    pos = numpy.nonzero(flag)[0].astype(numpy.int32) + 1
    return pos, data[pos]


def next_highest_power_of_2(value):
    p = 1
    while p < value:
        p <<= 1
    return p


def zeropadding(src, additional=0):
    """Copy into ``2 * pow2ceil(len) + additional`` floats, tail zero (``memory.h:120-170``)."""
    src = numpy.asarray(src, dtype=numpy.float32)
    out = numpy.zeros(2 * next_highest_power_of_2(src.size) + additional, dtype=numpy.float32)
    out[:src.size] = src
    return out


def complex_multiply(a, b, conjugate=False):
    """Interleaved (re, im) float arrays."""
    a = numpy.asarray(a, dtype=numpy.float32)
    b = numpy.asarray(b, dtype=numpy.float32)
    ca = a[..., 0::2] + 1j * a[..., 1::2]
    cb = b[..., 0::2] + 1j * b[..., 1::2]
    r = ca * (numpy.conj(cb) if conjugate else cb)
    out = numpy.empty(numpy.broadcast(a, b).shape, dtype=numpy.float32)
    out[..., 0::2] = r.real
    out[..., 1::2] = r.imag
    return out
