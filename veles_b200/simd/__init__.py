"""Signal / array primitives: the framework's counterpart of the reference's libSimd.

Reference: ``libVeles/simd/inc/simd/{memory,arithmetic,mathfun,matrix,convolve,correlate,normalize,wavelet,
detect_peaks}.h`` - a C library of SSE/AVX/NEON loops working on one signal per call.  Three implementations share
one API here, picked by what the caller passes:

* **CUDA tensors** (``torch`` float32 / uint8 on the B200) run the batched kernels of ``csrc/signal.cu``: every call
  takes ``[batch, length]`` rows, so a minibatch of audio frames or image planes is one launch and stays in HBM
  between feature-extraction stages.  Long filters go through ``torch.fft`` with the spectrum product in
  ``vb_sig_complex_mul``.
* **numpy arrays** run ``native/lib/libveles_simd.so`` (``native/simd/*.cc``, C++), when it has been built;
* otherwise the numpy oracle (``simd/oracle.py``), which is also what the tests compare the other two against.

The wavelet filter banks are computed by ``scripts/gen_wavelet_tables.py``; ``reference_scale=True`` (default)
reproduces the reference's table scaling (coiflets / symlets sum to 1, Daubechies to sqrt 2).
"""
from __future__ import annotations

import numpy

from . import oracle
from . import _native
from .oracle import (WAVELET_DAUBECHIES, WAVELET_COIFLET, WAVELET_SYMLET, EXT_PERIODIC, EXT_MIRROR, EXT_CONSTANT,  # noqa
                     EXT_ZERO, EXTREMUM_MAXIMUM, EXTREMUM_MINIMUM, EXTREMUM_BOTH, CONV_BRUTE, CONV_FFT,
                     CONV_OVERLAP_SAVE, wavelet_validate_order, wavelet_filters, convolve_choose,
                     next_highest_power_of_2, zeropadding)

#: device convolution switches from the direct kernel to FFT above this many taps
DEVICE_DIRECT_MAX_TAPS = 1024


def _is_cuda(x):
    try:
        import torch
    except ImportError:       # pragma: no cover
        return False
    return isinstance(x, torch.Tensor) and x.is_cuda


def _f32(a):
    return numpy.ascontiguousarray(a, dtype=numpy.float32)


def _rows2d(t):
    """A CUDA tensor as contiguous float32 rows + a function restoring the caller's leading shape."""
    import torch
    lead = t.shape[:-1]
    r = t.reshape(-1, t.shape[-1]).to(torch.float32)
    if r.stride(-1) != 1:
        r = r.contiguous()
    return r, (lambda out: out.reshape(lead + out.shape[1:]))


# ---- convolution / correlation -----------------------------------------------------------------------------------
def _device_fft_conv(x, h, reverse, overlap_save):
    import torch
    from ..ops import kernels
    B, X = x.shape
    M = h.shape[-1]
    out_len = X + M - 1
    hh = torch.flip(h, dims=(-1,)) if reverse else h
    if not overlap_save:
        n = next_highest_power_of_2(out_len)
        xs = torch.view_as_real(torch.fft.rfft(x, n=n)).reshape(B, -1)
        hs = torch.view_as_real(torch.fft.rfft(hh, n=n)).reshape(hh.shape[0], -1)
        prod = torch.empty_like(xs)
        kernels.sig_complex_mul(xs, hs, prod)
        y = torch.fft.irfft(torch.view_as_complex(prod.reshape(B, -1, 2)), n=n)
        return y[:, :out_len].contiguous()
    # overlap-save: blocks of L = 4 * 2^floor(log2 M) samples advance by L - (M - 1); each block's alias-free tail
    # is a run of outputs (reference block size: convolve.c:115-121)
    L = 1 << (2 + int(numpy.floor(numpy.log2(M))))
    step = L - (M - 1)
    nblk = (out_len + step - 1) // step
    padded = torch.zeros(B, (M - 1) + nblk * step + L, dtype=torch.float32, device=x.device)
    padded[:, M - 1:M - 1 + X] = x
    frames = padded.unfold(1, L, step)[:, :nblk]                       # [B, nblk, L] view
    spec = torch.view_as_real(torch.fft.rfft(frames, n=L)).reshape(B * nblk, -1)
    hs = torch.view_as_real(torch.fft.rfft(hh, n=L)).reshape(hh.shape[0], -1)
    if hs.shape[0] != 1:
        hs = hs.repeat_interleave(nblk, dim=0)
    prod = torch.empty_like(spec)
    kernels.sig_complex_mul(spec, hs, prod)
    y = torch.fft.irfft(torch.view_as_complex(prod.reshape(B * nblk, -1, 2)), n=L).reshape(B, nblk, L)
    return y[:, :, M - 1:].reshape(B, -1)[:, :out_len].contiguous()


def _convolve(x, h, algorithm, reverse):
    if _is_cuda(x):
        import torch
        from ..ops import kernels
        xr, restore = _rows2d(x)
        hr = h if _is_cuda(h) else torch.as_tensor(numpy.asarray(h, dtype=numpy.float32), device=x.device)
        hr = hr.reshape(-1, hr.shape[-1]).to(torch.float32).contiguous()
        M = hr.shape[-1]
        if algorithm is None:
            algorithm = CONV_BRUTE if M <= DEVICE_DIRECT_MAX_TAPS else (
                CONV_OVERLAP_SAVE if xr.shape[-1] > 4 * M else CONV_FFT)
        if algorithm == CONV_OVERLAP_SAVE and not M < xr.shape[-1] // 2:
            algorithm = CONV_FFT
        if algorithm == CONV_BRUTE:
            out = torch.empty(xr.shape[0], xr.shape[1] + M - 1, dtype=torch.float32, device=x.device)
            kernels.sig_convolve(xr, hr, out, reverse=reverse)
        else:
            out = _device_fft_conv(xr, hr, reverse, algorithm == CONV_OVERLAP_SAVE)
        return restore(out)
    x = _f32(x)
    h = _f32(h)
    L = _native.lib()
    if L is None:
        return oracle.cross_correlate(x, h) if reverse else oracle.convolve(x, h)
    if x.ndim == 1:
        init = L.vs_cross_correlate_initialize_algorithm if reverse else L.vs_convolve_initialize_algorithm
        alg = convolve_choose(x.size, h.size) if algorithm is None else algorithm
        handle = init(x.size, h.size, alg)
        out = numpy.empty(x.size + h.size - 1, dtype=numpy.float32)
        try:
            (L.vs_cross_correlate if reverse else L.vs_convolve)(handle, _native.ptr(x), _native.ptr(h),
                                                                _native.ptr(out))
        finally:
            L.vs_convolve_finalize(handle)
        return out
    hh = numpy.broadcast_to(h, (x.shape[0], h.shape[-1])) if h.ndim == 1 else h
    return numpy.stack([_convolve(a, _f32(b), algorithm, reverse) for a, b in zip(x, hh)])


def convolve(x, h, algorithm=None):
    """Full linear convolution (``len(x) + len(h) - 1`` outputs per row).  ``algorithm``: None = choose,
    ``CONV_BRUTE`` / ``CONV_FFT`` / ``CONV_OVERLAP_SAVE`` (reference: ``convolve.h:57-125``)."""
    return _convolve(x, h, algorithm, False)


def cross_correlate(x, h, algorithm=None):
    """``numpy.correlate(x, h, "full")`` per row (reference: ``correlate.h:58-135``)."""
    return _convolve(x, h, algorithm, True)


# ---- wavelets ------------------------------------------------------------------------------------------------------
def wavelet_apply(wtype, order, ext, src, reference_scale=True):
    """One decimated analysis step -> (highpass half, lowpass half) (reference: ``wavelet.h:96-104``)."""
    if _is_cuda(src):
        import torch
        from ..ops import kernels
        lo_t, hi_t = wavelet_filters(wtype, order, reference_scale)
        xr, restore = _rows2d(src)
        hi = torch.empty(xr.shape[0], xr.shape[1] // 2, dtype=torch.float32, device=src.device)
        lo = torch.empty_like(hi)
        kernels.sig_wavelet(xr, lo_t, hi_t, hi, lo, ext)
        return restore(hi), restore(lo)
    src = _f32(src)
    L = _native.lib()
    if L is None:
        return oracle.wavelet_apply(wtype, order, ext, src, reference_scale)
    if not wavelet_validate_order(wtype, order):
        raise ValueError("wavelet %r with %d taps is not available" % (wtype, order))
    if src.shape[-1] % 2:
        raise ValueError("the decimated transform needs an even length")
    rows = src.reshape(-1, src.shape[-1])
    hi = numpy.empty((rows.shape[0], rows.shape[1] // 2), dtype=numpy.float32)
    lo = numpy.empty_like(hi)
    tid = oracle.wavelet_type_id(wtype)
    for i in range(rows.shape[0]):
        L.vs_wavelet_apply_ex(tid, order, ext, int(reference_scale), _native.ptr(rows[i]), rows.shape[1],
                              _native.ptr(hi[i]), _native.ptr(lo[i]))
    shape = src.shape[:-1] + (rows.shape[1] // 2,)
    return hi.reshape(shape), lo.reshape(shape)


def stationary_wavelet_apply(wtype, order, level, ext, src, reference_scale=True):
    """One undecimated step at ``level`` -> (highpass, lowpass), input length (reference: ``wavelet.h:128-137``)."""
    if _is_cuda(src):
        import torch
        from ..ops import kernels
        lo_t, hi_t = wavelet_filters(wtype, order, reference_scale)
        xr, restore = _rows2d(src)
        hi = torch.empty_like(xr)
        lo = torch.empty_like(xr)
        kernels.sig_wavelet(xr, lo_t, hi_t, hi, lo, ext, level=level, stationary=True)
        return restore(hi), restore(lo)
    src = _f32(src)
    L = _native.lib()
    if L is None:
        return oracle.stationary_wavelet_apply(wtype, order, level, ext, src, reference_scale)
    if not wavelet_validate_order(wtype, order):
        raise ValueError("wavelet %r with %d taps is not available" % (wtype, order))
    rows = src.reshape(-1, src.shape[-1])
    hi = numpy.empty_like(rows)
    lo = numpy.empty_like(rows)
    tid = oracle.wavelet_type_id(wtype)
    for i in range(rows.shape[0]):
        L.vs_stationary_wavelet_apply_ex(tid, order, level, ext, int(reference_scale), _native.ptr(rows[i]),
                                         rows.shape[1], _native.ptr(hi[i]), _native.ptr(lo[i]))
    return hi.reshape(src.shape), lo.reshape(src.shape)


def wavelet_decompose(wtype, order, ext, src, levels, reference_scale=True):
    """Multi-level pyramid: the lowpass half is split again ``levels`` times -> [hi_1, hi_2, ..., hi_n, lo_n]."""
    out = []
    cur = src
    for _ in range(levels):
        hi, cur = wavelet_apply(wtype, order, ext, cur, reference_scale)
        out.append(hi)
    out.append(cur)
    return out


# ---- min / max / normalisation -------------------------------------------------------------------------------------
def minmax1D(src):
    """(min, max) of every row (reference: ``normalize.h:84-89``)."""
    if _is_cuda(src):
        import torch
        from ..ops import kernels
        xr, restore = _rows2d(src)
        out = torch.empty(xr.shape[0], 2, dtype=torch.float32, device=src.device)
        kernels.sig_minmax(xr, out)
        lead = src.shape[:-1]
        return out[:, 0].reshape(lead), out[:, 1].reshape(lead)
    src = _f32(src)
    L = _native.lib()
    if L is None or src.ndim != 1:
        return oracle.minmax1D(src)
    import ctypes
    lo, hi = ctypes.c_float(), ctypes.c_float()
    L.vs_minmax1D(_native.ptr(src), src.size, ctypes.byref(lo), ctypes.byref(hi))
    return numpy.float32(lo.value), numpy.float32(hi.value)


def _planes(t):
    import torch
    if t.dtype != torch.uint8:
        raise TypeError("8-bit planes expected")
    p = t.reshape((-1,) + tuple(t.shape[-2:]))
    return p if p.stride(-1) == 1 else p.contiguous()


def minmax2D(src):
    """(min, max) of every uint8 plane ``[..., H, W]`` (reference: ``normalize.h:63-66``)."""
    if _is_cuda(src):
        import torch
        from ..ops import kernels
        p = _planes(src)
        mm = torch.empty(p.shape[0], 2, dtype=torch.int32, device=src.device)
        kernels.sig_normalize_u8(p, mm, None, have_minmax=False)
        lead = src.shape[:-2]
        return mm[:, 0].to(torch.uint8).reshape(lead), mm[:, 1].to(torch.uint8).reshape(lead)
    src = numpy.ascontiguousarray(src, dtype=numpy.uint8)
    L = _native.lib()
    if L is None or src.ndim != 2:
        return oracle.minmax2D(src)
    import ctypes
    lo, hi = ctypes.c_uint8(), ctypes.c_uint8()
    L.vs_minmax2D(_native.ptr(src), src.shape[1], src.shape[1], src.shape[0], ctypes.byref(lo), ctypes.byref(hi))
    return numpy.uint8(lo.value), numpy.uint8(hi.value)


def normalize2D(src, minmax=None):
    """uint8 planes -> float32 in [-1, 1] by each plane's own range; ``minmax=(min, max)`` skips the range pass
    (reference: ``normalize2D`` / ``normalize2D_minmax``, ``normalize.h:48-82``)."""
    if _is_cuda(src):
        import torch
        from ..ops import kernels
        p = _planes(src)
        mm = torch.empty(p.shape[0], 2, dtype=torch.int32, device=src.device)
        if minmax is not None:
            mm[:, 0] = torch.as_tensor(minmax[0], device=src.device).reshape(-1).to(torch.int32)
            mm[:, 1] = torch.as_tensor(minmax[1], device=src.device).reshape(-1).to(torch.int32)
        dst = torch.empty(p.shape, dtype=torch.float32, device=src.device)
        kernels.sig_normalize_u8(p, mm, dst, have_minmax=minmax is not None)
        return dst.reshape(src.shape)
    src = numpy.ascontiguousarray(src, dtype=numpy.uint8)
    L = _native.lib()
    if L is None or src.ndim != 2:
        return oracle.normalize2D(src) if minmax is None else oracle.normalize2D_minmax(minmax[0], minmax[1], src)
    dst = numpy.empty(src.shape, dtype=numpy.float32)
    h, w = src.shape
    if minmax is None:
        L.vs_normalize2D(_native.ptr(src), w, w, h, _native.ptr(dst), w)
    else:
        L.vs_normalize2D_minmax(int(minmax[0]), int(minmax[1]), _native.ptr(src), w, w, h, _native.ptr(dst), w)
    return dst


def normalize2D_minmax(vmin, vmax, src):
    return normalize2D(src, minmax=(vmin, vmax))


# ---- peaks ---------------------------------------------------------------------------------------------------------
def detect_peaks(data, extremum_type=EXTREMUM_BOTH, capacity=None):
    """Strict interior extrema in index order (reference: ``detect_peaks.h``).

    One signal -> ``(positions int32, values float32)``.  A batch of rows -> ``(positions [B, cap], values [B, cap],
    counts [B])`` with ``cap`` = ``capacity`` or the worst case (``length - 2`` for both kinds, ``length // 2`` for
    one); entries past a row's count are undefined and a row's count may exceed a smaller ``capacity``.  On the device nothing is copied back: the three tensors stay on the GPU."""
    if _is_cuda(data):
        import torch
        from ..ops import kernels
        xr, _ = _rows2d(data)
        B, n = xr.shape
        # worst case: an alternating signal makes every interior sample an extremum (one kind: every other one)
        cap = int(capacity) if capacity else max(1, n - 2 if extremum_type == EXTREMUM_BOTH else n // 2)
        blocks = (n + 1023) // 1024
        counts = torch.empty(B * blocks, dtype=torch.int32, device=data.device)
        pos = torch.empty(B, cap, dtype=torch.int32, device=data.device)
        val = torch.empty(B, cap, dtype=torch.float32, device=data.device)
        totals = torch.empty(B, dtype=torch.int32, device=data.device)
        kernels.sig_detect_peaks(xr, extremum_type, counts, pos, val, totals)
        if data.dim() == 1:
            k = min(int(totals[0].item()), cap)
            return pos[0, :k], val[0, :k]
        return pos, val, totals
    data = _f32(data)
    if data.ndim == 2:
        found = [detect_peaks(r, extremum_type) for r in data]
        cap = int(capacity) if capacity else max([1] + [len(p) for p, _ in found])
        pos = numpy.zeros((len(found), cap), numpy.int32)
        val = numpy.zeros((len(found), cap), numpy.float32)
        for i, (p, v) in enumerate(found):
            k = min(len(p), cap)
            pos[i, :k], val[i, :k] = p[:k], v[:k]
        return pos, val, numpy.array([len(p) for p, _ in found], numpy.int32)
    L = _native.lib()
    if L is None:
        return oracle.detect_peaks(data, extremum_type)
    import ctypes
    res = ctypes.POINTER(_native.ExtremumPoint)()
    cnt = ctypes.c_size_t()
    L.vs_detect_peaks(_native.ptr(data), data.size, int(extremum_type), ctypes.byref(res), ctypes.byref(cnt))
    n = cnt.value
    pos = numpy.fromiter((res[i].position for i in range(n)), numpy.int32, n)
    val = numpy.fromiter((res[i].value for i in range(n)), numpy.float32, n)
    if n:
        L.vs_free(ctypes.cast(res, ctypes.c_void_p))
    return pos, val


# ---- arithmetic / conversions --------------------------------------------------------------------------------------
_CONVERSIONS = {  # name: (src numpy dtype, dst numpy dtype, device kind or None)
    "int16_to_float": (numpy.int16, numpy.float32, 0), "float_to_int16": (numpy.float32, numpy.int16, 1),
    "int32_to_float": (numpy.int32, numpy.float32, 2), "float_to_int32": (numpy.float32, numpy.int32, 3),
    "float16_to_float": (numpy.float16, numpy.float32, 4), "int16_to_int32": (numpy.int16, numpy.int32, None),
    "int32_to_int16": (numpy.int32, numpy.int16, None),
}


def _make_conversion(name):
    sdt, ddt, kind = _CONVERSIONS[name]

    def convert(data):
        if _is_cuda(data):
            import torch
            from ..ops import kernels
            tdt = {numpy.float32: torch.float32, numpy.int16: torch.int16, numpy.int32: torch.int32,
                   numpy.float16: torch.float16}
            if kind is None:
                return data.to(tdt[ddt])
            src = data.to(tdt[sdt]).contiguous()
            dst = torch.empty(src.shape, dtype=tdt[ddt], device=data.device)
            return kernels.sig_convert(src, dst, kind)
        src = numpy.ascontiguousarray(data, dtype=sdt)
        L = _native.lib()
        if L is None:
            with numpy.errstate(invalid="ignore"):
                return src.astype(ddt)              # numpy truncates float -> int like the reference does
        dst = numpy.empty(src.shape, dtype=ddt)
        getattr(L, "vs_" + name)(_native.ptr(src), src.size, _native.ptr(dst))
        return dst
    convert.__name__ = name
    convert.__doc__ = "Sample-format conversion (reference: ``arithmetic.h`` ``%s``)." % name
    return convert


for _n in _CONVERSIONS:
    globals()[_n] = _make_conversion(_n)


def real_multiply(a, b):
    if _is_cuda(a):
        return a * b
    a, b = _f32(a), _f32(b)
    L = _native.lib()
    if L is None or a.shape != b.shape:
        return a * b
    out = numpy.empty_like(a)
    L.vs_real_multiply_array(_native.ptr(a), _native.ptr(b), a.size, _native.ptr(out))
    return out


def real_multiply_scalar(a, value):
    if _is_cuda(a):
        return a * float(value)
    a = _f32(a)
    L = _native.lib()
    if L is None:
        return a * numpy.float32(value)
    out = numpy.empty_like(a)
    L.vs_real_multiply_scalar(_native.ptr(a), a.size, float(value), _native.ptr(out))
    return out


def complex_multiply(a, b, conjugate=False):
    """Interleaved (re, im) float32 arrays: ``a * b`` or ``a * conj(b)`` (reference: ``arithmetic.h:661-700``)."""
    if _is_cuda(a):
        import torch
        from ..ops import kernels
        ar, restore = _rows2d(a)
        br, _ = _rows2d(b)
        out = torch.empty_like(ar)
        kernels.sig_complex_mul(ar, br, out, conj_b=conjugate)
        return restore(out)
    a, b = _f32(a), _f32(b)
    L = _native.lib()
    if L is None or a.shape != b.shape:
        return oracle.complex_multiply(a, b, conjugate)
    out = numpy.empty_like(a)
    (L.vs_complex_multiply_conjugate if conjugate else L.vs_complex_multiply)(_native.ptr(a), _native.ptr(b), a.size,
                                                                             _native.ptr(out))
    return out


def complex_multiply_conjugate(a, b):
    return complex_multiply(a, b, conjugate=True)


def complex_conjugate(a):
    if _is_cuda(a):
        out = a.clone()
        out[..., 1::2] = -out[..., 1::2]
        return out
    a = _f32(a)
    out = a.copy()
    out[..., 1::2] *= -1
    return out


def sum_elements(a):
    if _is_cuda(a):
        return a.sum()
    a = _f32(a)
    L = _native.lib()
    return a.sum(dtype=numpy.float32) if L is None else numpy.float32(L.vs_sum_elements(_native.ptr(a), a.size))


def add_to_all(a, value):
    return a + float(value) if _is_cuda(a) else _f32(a) + numpy.float32(value)


def _psv(name, npf):
    def fn(src):
        if _is_cuda(src):
            import torch
            return getattr(torch, name)(src)
        src = _f32(src)
        L = _native.lib()
        if L is None:
            return npf(src)
        out = numpy.empty_like(src)
        getattr(L, "vs_%s_psv" % name)(_native.ptr(src), src.size, _native.ptr(out))
        return out
    fn.__name__ = name + "_psv"
    fn.__doc__ = "Element-wise ``%s`` (reference: ``mathfun.h:142-205``)." % name
    return fn


sin_psv, cos_psv, log_psv, exp_psv = (_psv("sin", numpy.sin), _psv("cos", numpy.cos), _psv("log", numpy.log),
                                      _psv("exp", numpy.exp))


# ---- matrices (row-major) ------------------------------------------------------------------------------------------
def matrix_add(m1, m2):
    return m1 + m2 if _is_cuda(m1) else _f32(m1) + _f32(m2)


def matrix_sub(m1, m2):
    return m1 - m2 if _is_cuda(m1) else _f32(m1) - _f32(m2)


def matrix_multiply(m1, m2, transposed=False):
    """``m1 @ m2`` (or ``m1 @ m2.T``) in float32 (reference: ``matrix.h:64-87``).  CUDA tensors use the tcgen05 TF32
    GEMM when the pitches satisfy TMA alignment (multiples of 4 floats), else ``torch.matmul``."""
    if _is_cuda(m1):
        import torch
        a = m1.to(torch.float32).contiguous()
        b = m2.to(torch.float32).contiguous()
        n = b.shape[0] if transposed else b.shape[1]
        if a.shape[1] % 4 == 0 and n % 4 == 0 and b.shape[1] % 4 == 0:
            from ..ops import kernels
            out = torch.empty(a.shape[0], n, dtype=torch.float32, device=a.device)
            kernels.gemm_tf32(a, b, out, b_mn=not transposed)
            return out
        return a @ (b.t() if transposed else b)
    a, b = _f32(m1), _f32(m2)
    L = _native.lib()
    if L is None:
        return a @ (b.T if transposed else b)
    if transposed:
        out = numpy.empty((a.shape[0], b.shape[0]), numpy.float32)
        L.vs_matrix_multiply_transposed(_native.ptr(a), _native.ptr(b), a.shape[1], a.shape[0], b.shape[1], b.shape[0],
                                        _native.ptr(out))
    else:
        out = numpy.empty((a.shape[0], b.shape[1]), numpy.float32)
        L.vs_matrix_multiply(_native.ptr(a), _native.ptr(b), a.shape[1], a.shape[0], b.shape[1], b.shape[0],
                             _native.ptr(out))
    return out


def matrix_multiply_transposed(m1, m2):
    return matrix_multiply(m1, m2, transposed=True)


def native_available():
    return _native.lib() is not None
