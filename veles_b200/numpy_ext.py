"""numpy helpers (reference: veles/numpy_ext.py:42-132)."""
import numpy


def roundup(num, align):
    d = num % align
    return num if d == 0 else num + (align - d)


def ravel(a):
    """Flatten without copying (asserts the memory address is preserved)."""
    b = a.ravel()
    assert b.__array_interface__["data"][0] == a.__array_interface__["data"][0], "ravel() copied"
    return b


def reshape(a, shape):
    b = a.reshape(shape)
    assert b.__array_interface__["data"][0] == a.__array_interface__["data"][0], "reshape() copied"
    return b


def transpose(a):
    b = a.transpose()
    assert b.__array_interface__["data"][0] == a.__array_interface__["data"][0], "transpose() copied"
    return b


def interleave(arr):
    """NCHW -> NHWC (or CHW -> HWC)."""
    if arr.ndim == 4:
        return numpy.ascontiguousarray(arr.transpose(0, 2, 3, 1))
    if arr.ndim == 3:
        return numpy.ascontiguousarray(arr.transpose(1, 2, 0))
    raise ValueError("interleave() expects a 3- or 4-dimensional array")


def norm_image(a, yuv=False):
    aa = a.astype(numpy.float32)
    lo, hi = aa.min(), aa.max()
    if hi > lo:
        aa = (aa - lo) * (255.0 / (hi - lo))
    else:
        aa[:] = 127.5
    return aa.astype(numpy.uint8)


class NumDiff(object):
    """5-point numeric derivative helper for gradient checks (reference :116-132)."""

    def __init__(self):
        self.h = 1.0e-3
        self.points = (2.0 * self.h, self.h, -self.h, -2.0 * self.h)
        self.coeffs = numpy.array([-1.0, 8.0, -8.0, 1.0], dtype=numpy.float64)
        self.divizor = 12.0 * self.h
        self.errs = numpy.zeros_like(self.points)

    def check_diff(self, x, i, fn):
        """d fn / d x.flat[i] where fn() evaluates the scalar with the current x."""
        flat = x.reshape(-1)
        saved = flat[i]
        for k, p in enumerate(self.points):
            flat[i] = saved + p
            self.errs[k] = fn()
        flat[i] = saved
        return float((self.errs * self.coeffs).sum() / self.divizor)

    @property
    def derivative(self):
        return (self.errs * self.coeffs).sum() / self.divizor
