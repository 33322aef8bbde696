"""ctypes binding of ``native/lib/libveles_simd.so`` (``native/include/veles/simd.h``)."""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_float, c_int, c_size_t, c_void_p

_LIB = None
_TRIED = False


class ExtremumPoint(ctypes.Structure):
    _fields_ = [("position", c_int), ("value", c_float)]


def path():
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    return os.path.join(root, "native", "lib", "libveles_simd.so")


def lib(required=False):
    """The loaded library or None when it has not been built (``python native/build.py``)."""
    global _LIB, _TRIED
    if _LIB is None and not _TRIED:
        _TRIED = True
        p = path()
        if os.path.exists(p):
            L = ctypes.CDLL(p)
            fp, sz = POINTER(c_float), c_size_t
            L.vs_malloc_aligned.restype = c_void_p
            L.vs_malloc_aligned.argtypes = [sz]
            L.vs_free.argtypes = [c_void_p]
            L.vs_next_highest_power_of_2.restype = sz
            L.vs_next_highest_power_of_2.argtypes = [sz]
            L.vs_sum_elements.restype = c_float
            L.vs_convolve_initialize.restype = c_void_p
            L.vs_convolve_initialize.argtypes = [sz, sz]
            L.vs_convolve_initialize_algorithm.restype = c_void_p
            L.vs_convolve_initialize_algorithm.argtypes = [sz, sz, c_int]
            L.vs_cross_correlate_initialize.restype = c_void_p
            L.vs_cross_correlate_initialize.argtypes = [sz, sz]
            L.vs_cross_correlate_initialize_algorithm.restype = c_void_p
            L.vs_cross_correlate_initialize_algorithm.argtypes = [sz, sz, c_int]
            L.vs_convolve_algorithm.argtypes = [c_void_p]
            L.vs_convolve.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p]
            L.vs_cross_correlate.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p]
            L.vs_convolve_finalize.argtypes = [c_void_p]
            L.vs_convolve_simd.argtypes = [c_void_p, sz, c_void_p, sz, c_void_p]
            L.vs_cross_correlate_simd.argtypes = [c_void_p, sz, c_void_p, sz, c_void_p]
            L.vs_convolve_choose.argtypes = [sz, sz]
            L.vs_fft_complex.argtypes = [c_void_p, sz, c_int]
            L.vs_wavelet_filters.argtypes = [c_int, c_int, c_int, c_void_p, c_void_p]
            L.vs_wavelet_apply_ex.argtypes = [c_int, c_int, c_int, c_int, c_void_p, sz, c_void_p, c_void_p]
            L.vs_stationary_wavelet_apply_ex.argtypes = [c_int, c_int, c_int, c_int, c_int, c_void_p, sz, c_void_p,
                                                         c_void_p]
            L.vs_detect_peaks.argtypes = [c_void_p, sz, c_int, POINTER(POINTER(ExtremumPoint)), POINTER(sz)]
            L.vs_minmax1D.argtypes = [c_void_p, c_int, fp, fp]
            L.vs_minmax2D.argtypes = [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]
            L.vs_normalize2D.argtypes = [c_void_p, c_int, c_int, c_int, c_void_p, c_int]
            L.vs_normalize2D_minmax.argtypes = [ctypes.c_uint8, ctypes.c_uint8, c_void_p, c_int, c_int, c_int, c_void_p,
                                                c_int]
            for name in ("vs_matrix_add", "vs_matrix_sub"):
                getattr(L, name).argtypes = [c_void_p, c_void_p, sz, sz, c_void_p]
            for name in ("vs_matrix_multiply", "vs_matrix_multiply_transposed"):
                getattr(L, name).argtypes = [c_void_p, c_void_p, sz, sz, sz, sz, c_void_p]
            for name in ("vs_int16_to_float", "vs_float_to_int16", "vs_int32_to_float", "vs_float_to_int32",
                         "vs_int16_to_int32", "vs_int32_to_int16", "vs_float16_to_float", "vs_complex_conjugate",
                         "vs_sin_psv", "vs_cos_psv", "vs_log_psv", "vs_exp_psv"):
                getattr(L, name).argtypes = [c_void_p, sz, c_void_p]
            for name in ("vs_real_multiply_array", "vs_complex_multiply", "vs_complex_multiply_conjugate",
                         "vs_int16_multiply"):
                getattr(L, name).argtypes = [c_void_p, c_void_p, sz, c_void_p]
            L.vs_real_multiply_scalar.argtypes = [c_void_p, sz, c_float, c_void_p]
            L.vs_sum_elements.argtypes = [c_void_p, sz]
            L.vs_add_to_all.argtypes = [c_void_p, sz, c_float]
            L.vs_memsetf.argtypes = [c_void_p, c_float, sz]
            L.vs_rmemcpyf.argtypes = [c_void_p, c_void_p, sz]
            L.vs_rmemcpyf.restype = c_void_p
            L.vs_crmemcpyf.argtypes = [c_void_p, c_void_p, sz]
            L.vs_crmemcpyf.restype = c_void_p
            _LIB = L
    if _LIB is None and required:
        raise RuntimeError("native signal library missing: run `python native/build.py` (%s)" % path())
    return _LIB


def ptr(a):
    return a.ctypes.data_as(c_void_p)
