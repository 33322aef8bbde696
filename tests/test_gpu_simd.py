"""Batched signal kernels (csrc/signal.cu through veles_b200.simd) against the numpy oracle on a B200."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("CUDA device required", allow_module_level=True)

from veles_b200 import simd                      # noqa: E402
from veles_b200.simd import oracle               # noqa: E402
from veles_b200.ops import kernels               # noqa: E402

RNG = np.random.RandomState(99)
DEV = torch.device("cuda", 0)


def dev(a):
    return torch.as_tensor(a, device=DEV)


@pytest.mark.parametrize("xl,hl,batch", [(50, 7, 1), (1000, 33, 4), (1024, 1, 2), (3000, 513, 3), (5000, 1024, 2),
                                         (17, 40, 5), (2047, 2, 1)])
def test_direct_convolution_kernel(xl, hl, batch):
    x = RNG.randn(batch, xl).astype(np.float32)
    h = RNG.randn(hl).astype(np.float32)
    before = kernels.LAUNCHES[0]
    got = simd.convolve(dev(x), dev(h), simd.CONV_BRUTE).cpu().numpy()
    assert kernels.LAUNCHES[0] == before + 1
    want = oracle.convolve(x, h)
    tol = 3e-5 * max(1.0, np.abs(want).max())
    assert got.shape == want.shape and np.abs(got - want).max() < tol
    gotc = simd.cross_correlate(dev(x), dev(h), simd.CONV_BRUTE).cpu().numpy()
    assert np.abs(gotc - oracle.cross_correlate(x, h)).max() < tol
    # one filter per row
    hb = RNG.randn(batch, hl).astype(np.float32)
    got = simd.convolve(dev(x), dev(hb), simd.CONV_BRUTE).cpu().numpy()
    assert np.abs(got - oracle.convolve(x, hb)).max() < tol


@pytest.mark.parametrize("xl,hl,batch", [(4096, 129, 3), (10000, 2049, 2), (700, 500, 2)])
def test_fft_and_overlap_save_paths(xl, hl, batch):
    x = RNG.randn(batch, xl).astype(np.float32)
    h = RNG.randn(hl).astype(np.float32)
    want = oracle.convolve(x, h)
    wantc = oracle.cross_correlate(x, h)
    tol = 1e-4 * max(1.0, np.abs(want).max())
    for alg in (simd.CONV_FFT, simd.CONV_OVERLAP_SAVE, None):
        got = simd.convolve(dev(x), dev(h), alg).cpu().numpy()
        assert got.shape == want.shape and np.abs(got - want).max() < tol, alg
        gotc = simd.cross_correlate(dev(x), dev(h), alg).cpu().numpy()
        assert np.abs(gotc - wantc).max() < tol, alg


@pytest.mark.parametrize("family,order", [("daubechies", 2), ("daubechies", 8), ("coiflet", 18), ("symlet", 20),
                                           ("daubechies", 76)])
@pytest.mark.parametrize("ext", [simd.EXT_PERIODIC, simd.EXT_MIRROR, simd.EXT_CONSTANT, simd.EXT_ZERO])
def test_wavelet_kernels(family, order, ext):
    for n, batch in ((max(order, 8) + (max(order, 8) % 2), 3), (1000, 7), (4096, 2)):
        x = RNG.randn(batch, n).astype(np.float32)
        hi, lo = simd.wavelet_apply(family, order, ext, dev(x))
        rhi, rlo = oracle.wavelet_apply(family, order, ext, x)
        assert hi.shape == (batch, n // 2)
        assert np.allclose(hi.cpu().numpy(), rhi, atol=3e-5) and np.allclose(lo.cpu().numpy(), rlo, atol=3e-5)
        for level in (1, 3):
            shi, slo = simd.stationary_wavelet_apply(family, order, level, ext, dev(x))
            rshi, rslo = oracle.stationary_wavelet_apply(family, order, level, ext, x)
            assert np.allclose(shi.cpu().numpy(), rshi, atol=3e-5) and np.allclose(slo.cpu().numpy(), rslo, atol=3e-5)


def test_wavelet_pyramid_stays_on_device():
    x = dev(RNG.randn(16, 2048).astype(np.float32))
    pyr = simd.wavelet_decompose("daubechies", 8, simd.EXT_PERIODIC, x, 4)
    assert all(p.is_cuda for p in pyr) and [p.shape[1] for p in pyr] == [1024, 512, 256, 128, 128]
    energy = sum(float((p.double() ** 2).sum()) for p in pyr)
    assert abs(energy - float((x.double() ** 2).sum())) < 1e-3 * energy


def test_minmax_and_normalize():
    x = RNG.randn(9, 12345).astype(np.float32)
    lo, hi = simd.minmax1D(dev(x))
    assert np.array_equal(lo.cpu().numpy(), x.min(1)) and np.array_equal(hi.cpu().numpy(), x.max(1))
    img = RNG.randint(3, 250, size=(6, 61, 47)).astype(np.uint8)
    img[5] = 77                                              # flat plane -> zeros
    lo, hi = simd.minmax2D(dev(img))
    assert np.array_equal(lo.cpu().numpy(), img.min((1, 2))) and np.array_equal(hi.cpu().numpy(), img.max((1, 2)))
    out = simd.normalize2D(dev(img)).cpu().numpy()
    assert np.allclose(out, oracle.normalize2D(img), atol=1e-6) and np.all(out[5] == 0)
    out = simd.normalize2D_minmax(0, 255, dev(img)).cpu().numpy()
    assert np.allclose(out, img.astype(np.float32) / 127.5 - 1, atol=1e-6)
    # a strided view (cropped planes) is read in place
    crop = dev(img)[:, 5:40, 3:30]
    assert np.allclose(simd.normalize2D(crop).cpu().numpy(), oracle.normalize2D(img[:, 5:40, 3:30]), atol=1e-6)


@pytest.mark.parametrize("n", [3, 12, 1024, 1025, 5000, 70000])
def test_detect_peaks_kernel(n):
    x = RNG.randn(4, n).astype(np.float32)
    x[1] = np.round(x[1])                                      # plateaus: ties are not peaks
    for t in (simd.EXTREMUM_MAXIMUM, simd.EXTREMUM_MINIMUM, simd.EXTREMUM_BOTH):
        pos, val, cnt = simd.detect_peaks(dev(x), t)
        pos, val, cnt = pos.cpu().numpy(), val.cpu().numpy(), cnt.cpu().numpy()
        for r in range(4):
            rp, rv = oracle.detect_peaks(x[r], t)
            assert cnt[r] == rp.size
            assert np.array_equal(pos[r, :cnt[r]], rp) and np.array_equal(val[r, :cnt[r]], rv)
    p1, v1 = simd.detect_peaks(dev(x[0]), simd.EXTREMUM_BOTH)
    rp, rv = oracle.detect_peaks(x[0], simd.EXTREMUM_BOTH)
    assert np.array_equal(p1.cpu().numpy(), rp)
    # a capacity smaller than the number of peaks truncates, the total is still reported
    pos, val, cnt = simd.detect_peaks(dev(x), simd.EXTREMUM_BOTH, capacity=1)
    assert pos.shape == (4, 1) and int(cnt[0]) == rp.size


def test_conversions_and_products():
    i16 = RNG.randint(-32768, 32767, 100003).astype(np.int16)
    assert np.array_equal(simd.int16_to_float(dev(i16)).cpu().numpy(), i16.astype(np.float32))
    f = (RNG.randn(100003) * 1000).astype(np.float32)
    assert np.array_equal(simd.float_to_int16(dev(f)).cpu().numpy(), f.astype(np.int16))
    assert np.array_equal(simd.float_to_int32(dev(f)).cpu().numpy(), f.astype(np.int32))
    i32 = RNG.randint(-10 ** 6, 10 ** 6, 777).astype(np.int32)
    assert np.array_equal(simd.int32_to_float(dev(i32)).cpu().numpy(), i32.astype(np.float32))
    half = RNG.randn(999).astype(np.float16)
    assert np.array_equal(simd.float16_to_float(dev(half)).cpu().numpy(), half.astype(np.float32))
    a = RNG.randn(5, 256).astype(np.float32)
    b = RNG.randn(5, 256).astype(np.float32)
    assert np.allclose(simd.complex_multiply(dev(a), dev(b)).cpu().numpy(), oracle.complex_multiply(a, b), atol=1e-5)
    assert np.allclose(simd.complex_multiply_conjugate(dev(a), dev(b)).cpu().numpy(),
                       oracle.complex_multiply(a, b, True), atol=1e-5)
    assert np.allclose(simd.complex_multiply(dev(a), dev(b[:1])).cpu().numpy(), oracle.complex_multiply(a, b[:1]),
                       atol=1e-5)
    m1 = RNG.randn(64, 128).astype(np.float32)
    m2 = RNG.randn(128, 32).astype(np.float32)
    got = simd.matrix_multiply(dev(m1), dev(m2)).cpu().numpy()
    assert np.allclose(got, m1 @ m2, atol=5e-2)               # TF32 operands
    got = simd.matrix_multiply_transposed(dev(m1), dev(m2.T.copy())).cpu().numpy()
    assert np.allclose(got, m1 @ m2, atol=5e-2)
