"""Signal / array primitives (veles_b200.simd): native C++ vs numpy oracle vs independent statements (scipy / the
defining equations) on the CPU; the batched CUDA kernels vs the oracle on the GPU.
Reference behaviour: libVeles/simd/tests/*.cc (convolve, correlate, wavelet, normalize, detect_peaks, matrix, memory,
arithmetic)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from veles_b200 import simd
from veles_b200.simd import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RNG = np.random.RandomState(1234)


@pytest.fixture(scope="module", autouse=True)
def native_lib():
    """Build native/lib/libveles_simd.so once (g++ only, seconds) so that the numpy path below is the C++ one."""
    if not simd.native_available():
        import runpy
        ns = runpy.run_path(os.path.join(ROOT, "native", "build.py"))
        ns["build_simd"]()
        simd._native._TRIED = False
    assert simd.native_available()


# ---- filter banks ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("family,orders", [("daubechies", range(2, 78, 2)), ("symlet", range(2, 78, 2)),
                                            ("coiflet", range(6, 36, 6))])
def test_filter_banks_are_orthonormal(family, orders):
    for order in orders:
        assert simd.wavelet_validate_order(family, order)
        h = oracle.wavelet_tables()[family][order]
        assert h.size == order
        assert abs(h.sum() - np.sqrt(2)) < 1e-8
        assert abs(np.dot(h, h) - 1) < 1e-8
        for m in range(1, order // 2):
            assert abs(np.dot(h[:-2 * m], h[2 * m:])) < 1e-8, (family, order, m)


def test_vanishing_moments():
    """Daubechies / symlet with 2N taps kill polynomials of degree < N; a coiflet with 6K taps has 2K vanishing
    wavelet moments and 2K - 1 vanishing scaling moments around tap 2K."""
    for family in ("daubechies", "symlet"):
        for order in (4, 8, 12, 20):
            h = oracle.wavelet_tables()[family][order]
            k = np.arange(order)
            g = h[::-1] * np.where(k % 2 == 0, 1.0, -1.0)
            for p in range(order // 2):
                assert abs(np.dot(g, (k / order) ** p)) < 1e-7, (family, order, p)
    for order in (6, 12, 18, 24, 30):
        K = order // 6
        h = oracle.wavelet_tables()["coiflet"][order]
        k = np.arange(order)
        g = h[::-1] * np.where(k % 2 == 0, 1.0, -1.0)
        for p in range(2 * K):
            assert abs(np.dot(g, ((k - (order - 1 - 2 * K)) / order) ** p)) < 1e-7
        for p in range(1, 2 * K):
            assert abs(np.dot(h, ((k - 2 * K) / order) ** p)) < 1e-7


def test_known_closed_forms():
    s3 = np.sqrt(3.0)
    db2 = np.array([1 + s3, 3 + s3, 3 - s3, 1 - s3]) / (4 * np.sqrt(2))
    assert np.allclose(oracle.wavelet_tables()["daubechies"][4], db2, atol=1e-12)
    assert np.allclose(oracle.wavelet_tables()["daubechies"][2], [np.sqrt(0.5)] * 2)
    lo, hi = simd.wavelet_filters("symlet", 2)          # reference scaling: symlets / coiflets sum to 1
    assert np.allclose(lo, [0.5, 0.5]) and np.allclose(hi, [0.5, -0.5])
    lo, hi = simd.wavelet_filters("daubechies", 4)
    assert np.allclose(hi, [db2[3], -db2[2], db2[1], -db2[0]], atol=1e-7)
    assert not simd.wavelet_validate_order("daubechies", 3)
    assert not simd.wavelet_validate_order("coiflet", 8)


def test_generator_reproduces_shipped_tables():
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import gen_wavelet_tables as g
    for N in (2, 5, 9):
        assert np.allclose(g.daubechies(N), oracle.wavelet_tables()["daubechies"][2 * N], atol=1e-12)
        assert np.allclose(g.symlet(N), oracle.wavelet_tables()["symlet"][2 * N], atol=1e-12)
    assert np.allclose(g.coiflet(2), oracle.wavelet_tables()["coiflet"][12], atol=1e-12)    # K = 1, then continuation


# ---- transforms -----------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("family,order", [("daubechies", 2), ("daubechies", 8), ("daubechies", 16), ("coiflet", 12),
                                           ("symlet", 6), ("symlet", 20), ("daubechies", 76)])
@pytest.mark.parametrize("ext", [simd.EXT_PERIODIC, simd.EXT_MIRROR, simd.EXT_CONSTANT, simd.EXT_ZERO])
def test_wavelet_native_matches_oracle(family, order, ext):
    for n in (max(order, 8), 2 * order + 6, 512):
        n += n % 2
        x = RNG.randn(n).astype(np.float32)
        hi, lo = simd.wavelet_apply(family, order, ext, x)
        rhi, rlo = oracle.wavelet_apply(family, order, ext, x)
        assert hi.shape == (n // 2,)
        assert np.allclose(hi, rhi, atol=2e-5) and np.allclose(lo, rlo, atol=2e-5)
        for level in (1, 2, 3):
            shi, slo = simd.stationary_wavelet_apply(family, order, level, ext, x)
            rshi, rslo = oracle.stationary_wavelet_apply(family, order, level, ext, x)
            assert shi.shape == (n,)
            assert np.allclose(shi, rshi, atol=2e-5) and np.allclose(slo, rslo, atol=2e-5)


def test_periodic_dwt_preserves_energy_and_reconstructs():
    """An orthonormal bank with periodic extension is an orthogonal map: energy is kept and the synthesis bank
    (transposed analysis) gives the signal back."""
    n, order = 256, 8
    x = RNG.randn(n).astype(np.float32)
    hi, lo = simd.wavelet_apply("daubechies", order, simd.EXT_PERIODIC, x)
    assert abs(np.dot(hi, hi) + np.dot(lo, lo) - np.dot(x, x)) < 1e-3 * np.dot(x, x)
    lo_t, hi_t = simd.wavelet_filters("daubechies", order)
    rec = np.zeros(n, np.float64)
    for d in range(n // 2):
        for j in range(order):
            rec[(2 * d + j) % n] += lo_t[j] * lo[d] + hi_t[j] * hi[d]
    assert np.allclose(rec, x, atol=1e-4)
    # orthonormal option for the families whose reference tables sum to 1
    hi, lo = simd.wavelet_apply("symlet", 8, simd.EXT_PERIODIC, x, reference_scale=False)
    assert abs(np.dot(hi, hi) + np.dot(lo, lo) - np.dot(x, x)) < 1e-3 * np.dot(x, x)
    hi1, lo1 = simd.wavelet_apply("symlet", 8, simd.EXT_PERIODIC, x)
    assert np.allclose(hi1 * np.sqrt(2), hi, atol=1e-5)


def test_wavelet_batch_and_pyramid():
    x = RNG.randn(5, 128).astype(np.float32)
    hi, lo = simd.wavelet_apply("coiflet", 6, simd.EXT_MIRROR, x)
    assert hi.shape == (5, 64)
    for i in range(5):
        h1, l1 = oracle.wavelet_apply("coiflet", 6, simd.EXT_MIRROR, x[i])
        assert np.allclose(hi[i], h1, atol=1e-5) and np.allclose(lo[i], l1, atol=1e-5)
    pyr = simd.wavelet_decompose("daubechies", 4, simd.EXT_PERIODIC, x[0], 3)
    assert [p.size for p in pyr] == [64, 32, 16, 16]
    with pytest.raises(ValueError):
        simd.wavelet_apply("daubechies", 4, simd.EXT_ZERO, x[0][:63])
    with pytest.raises(ValueError):
        simd.wavelet_apply("coiflet", 10, simd.EXT_ZERO, x[0])


# ---- convolution ----------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("xl,hl", [(50, 7), (201, 9), (300, 20), (400, 300), (1000, 33), (4096, 129), (64, 64),
                                   (17, 40), (3, 1), (1, 5)])
def test_convolution_algorithms_agree_with_scipy(xl, hl):
    from scipy import signal
    x = RNG.randn(xl).astype(np.float32)
    h = RNG.randn(hl).astype(np.float32)
    want = signal.convolve(x.astype(np.float64), h.astype(np.float64))
    wantc = signal.correlate(x.astype(np.float64), h.astype(np.float64), mode="full")
    tol = 2e-5 * max(1.0, np.abs(want).max())
    for alg in (None, simd.CONV_BRUTE, simd.CONV_FFT, simd.CONV_OVERLAP_SAVE):
        got = simd.convolve(x, h, alg)
        assert got.shape == (xl + hl - 1,) and got.dtype == np.float32
        assert np.abs(got - want).max() < tol, (alg, np.abs(got - want).max())
        gotc = simd.cross_correlate(x, h, alg)
        assert np.abs(gotc - wantc).max() < tol, (alg, np.abs(gotc - wantc).max())


def test_convolution_heuristic_follows_the_reference():
    assert simd.convolve_choose(100, 10) == simd.CONV_BRUTE
    assert simd.convolve_choose(201, 10) == simd.CONV_OVERLAP_SAVE
    assert simd.convolve_choose(300, 200) == simd.CONV_BRUTE
    assert simd.convolve_choose(351, 200) == simd.CONV_FFT
    L = simd._native.lib()
    for xl, hl in [(100, 10), (201, 10), (300, 200), (351, 200)]:
        h = L.vs_convolve_initialize(xl, hl)
        assert L.vs_convolve_algorithm(h) == simd.convolve_choose(xl, hl)
        L.vs_convolve_finalize(h)


def test_fft_roundtrip():
    L = simd._native.lib()
    z = RNG.randn(2 * 64).astype(np.float32)
    want = np.fft.fft(z[0::2] + 1j * z[1::2])
    buf = z.copy()
    assert L.vs_fft_complex(simd._native.ptr(buf), 64, 1) == 0
    assert np.allclose(buf[0::2] + 1j * buf[1::2], want, atol=1e-4)
    assert L.vs_fft_complex(simd._native.ptr(buf), 64, -1) == 0
    assert np.allclose(buf / 64, z, atol=1e-5)
    assert L.vs_fft_complex(simd._native.ptr(buf), 48, 1) == -1


# ---- normalisation / peaks ------------------------------------------------------------------------------------------
def test_normalize_and_minmax():
    img = RNG.randint(7, 200, size=(37, 53)).astype(np.uint8)
    lo, hi = simd.minmax2D(img)
    assert (lo, hi) == (img.min(), img.max())
    out = simd.normalize2D(img)
    assert out.dtype == np.float32 and abs(out.min() + 1) < 1e-6 and abs(out.max() - 1) < 1e-6
    assert np.allclose(out, oracle.normalize2D(img), atol=1e-6)
    assert np.allclose(simd.normalize2D_minmax(0, 255, img), (img.astype(np.float32) / 127.5) - 1, atol=1e-6)
    assert np.all(simd.normalize2D(np.full((4, 4), 9, np.uint8)) == 0)
    x = RNG.randn(1001).astype(np.float32)
    assert simd.minmax1D(x) == (x.min(), x.max())
    batch = RNG.randint(0, 255, size=(3, 8, 8)).astype(np.uint8)
    assert np.allclose(simd.normalize2D(batch), oracle.normalize2D(batch))


def test_detect_peaks():
    x = np.array([0, 1, 0, -1, 0, 0, 2, 2, 1, 3, -5, 4], np.float32)
    pos, val = simd.detect_peaks(x, simd.EXTREMUM_BOTH)
    assert pos.tolist() == [1, 3, 8, 9, 10] and val.tolist() == [1, -1, 1, 3, -5]
    assert simd.detect_peaks(x, simd.EXTREMUM_MAXIMUM)[0].tolist() == [1, 9]
    assert simd.detect_peaks(x, simd.EXTREMUM_MINIMUM)[0].tolist() == [3, 8, 10]
    y = RNG.randn(5000).astype(np.float32)
    for t in (1, 2, 3):
        p, v = simd.detect_peaks(y, t)
        rp, rv = oracle.detect_peaks(y, t)
        assert np.array_equal(p, rp) and np.array_equal(v, rv)
    assert simd.detect_peaks(np.arange(10, dtype=np.float32))[0].size == 0
    pos, val, cnt = simd.detect_peaks(np.stack([x, -x]))
    assert cnt.tolist() == [5, 5] and pos[1, :5].tolist() == [1, 3, 8, 9, 10]


# ---- arithmetic / matrices / memory ---------------------------------------------------------------------------------
def test_conversions_and_arithmetic():
    i16 = RNG.randint(-32768, 32767, 1003).astype(np.int16)
    assert np.array_equal(simd.int16_to_float(i16), i16.astype(np.float32))
    f = (RNG.randn(1003) * 1000).astype(np.float32)
    assert np.array_equal(simd.float_to_int16(f), f.astype(np.int16))          # truncation toward zero
    assert np.array_equal(simd.float_to_int32(f), f.astype(np.int32))
    i32 = RNG.randint(-10 ** 6, 10 ** 6, 77).astype(np.int32)
    assert np.array_equal(simd.int32_to_float(i32), i32.astype(np.float32))
    assert np.array_equal(simd.int16_to_int32(i16), i16.astype(np.int32))
    assert np.array_equal(simd.int32_to_int16(i16.astype(np.int32)), i16)
    half = np.concatenate([RNG.randn(500).astype(np.float16), np.array([0, -0.0, 6e-8, 65504, np.inf], np.float16)])
    assert np.array_equal(simd.float16_to_float(half), half.astype(np.float32))
    a, b = RNG.randn(64).astype(np.float32), RNG.randn(64).astype(np.float32)
    assert np.allclose(simd.real_multiply(a, b), a * b)
    assert np.allclose(simd.real_multiply_scalar(a, 2.5), a * 2.5)
    ca, cb = a[0::2] + 1j * a[1::2], b[0::2] + 1j * b[1::2]
    r = simd.complex_multiply(a, b)
    assert np.allclose(r[0::2] + 1j * r[1::2], ca * cb, atol=1e-6)
    r = simd.complex_multiply_conjugate(a, b)
    assert np.allclose(r[0::2] + 1j * r[1::2], ca * np.conj(cb), atol=1e-6)
    r = simd.complex_conjugate(a)
    assert np.allclose(r[0::2] + 1j * r[1::2], np.conj(ca))
    assert abs(simd.sum_elements(a) - a.sum()) < 1e-4
    assert np.allclose(simd.add_to_all(a, 3), a + 3)
    pos = np.abs(a) + 0.1
    for fn, ref in ((simd.sin_psv, np.sin), (simd.cos_psv, np.cos), (simd.log_psv, np.log), (simd.exp_psv, np.exp)):
        assert np.allclose(fn(pos), ref(pos), rtol=1e-5, atol=1e-6)


def test_matrices_and_memory():
    a = RNG.randn(17, 33).astype(np.float32)
    b = RNG.randn(33, 9).astype(np.float32)
    c = RNG.randn(21, 33).astype(np.float32)
    assert np.allclose(simd.matrix_multiply(a, b), a @ b, atol=1e-4)
    assert np.allclose(simd.matrix_multiply_transposed(a, c), a @ c.T, atol=1e-4)
    assert np.allclose(simd.matrix_add(a, a), 2 * a) and np.allclose(simd.matrix_sub(a, a), 0)
    assert simd.next_highest_power_of_2(100) == 128 and simd.next_highest_power_of_2(128) == 128
    z = simd.zeropadding(np.ones(100, np.float32))
    assert z.size == 256 and z[:100].sum() == 100 and z[100:].sum() == 0
    L = simd._native.lib()
    import ctypes
    p = L.vs_malloc_aligned(1000)
    assert p % 64 == 0
    L.vs_free(p)
    src = np.arange(10, dtype=np.float32)
    dst = np.empty_like(src)
    L.vs_rmemcpyf(simd._native.ptr(dst), simd._native.ptr(src), 10)
    assert np.array_equal(dst, src[::-1])
    L.vs_crmemcpyf(simd._native.ptr(dst), simd._native.ptr(src), 10)
    assert dst.tolist() == [8, 9, 6, 7, 4, 5, 2, 3, 0, 1]
    L.vs_memsetf(simd._native.ptr(dst), 1.5, 10)
    assert np.all(dst == 1.5)
    n = ctypes.c_size_t()
    L.vs_zeropaddingex.restype = ctypes.c_void_p
    L.vs_zeropaddingex.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t), ctypes.c_size_t]
    q = L.vs_zeropaddingex(simd._native.ptr(src), 10, ctypes.byref(n), 5)
    assert n.value == 2 * 16 + 5
    L.vs_free(q)


def test_native_test_binary():
    exe = os.path.join(ROOT, "native", "lib", "test_simd")
    if not os.path.exists(exe):
        pytest.skip("native/lib/test_simd not built")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
