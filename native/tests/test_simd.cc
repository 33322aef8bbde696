// Self-checking test binary for libveles_simd (no test framework: each CHECK prints and counts failures).
// Mirrors what the reference covers in libVeles/simd/tests/{convolve,correlate,wavelet,normalize,detect_peaks,
// matrix,memory,arithmetic}.cc: every accelerated path against a plain restatement written here.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "veles/simd.h"

static int g_failed = 0;
#define CHECK(cond)                                                   \
  do {                                                                \
    if (!(cond)) {                                                    \
      printf("FAILED %s:%d  %s\n", __FILE__, __LINE__, #cond);        \
      ++g_failed;                                                     \
    }                                                                 \
  } while (0)

static std::vector<float> noise(size_t n, unsigned seed) {
  std::vector<float> v(n);
  unsigned s = seed * 2654435761u + 1;
  for (size_t i = 0; i < n; ++i) {
    s = s * 1664525u + 1013904223u;
    v[i] = static_cast<float>(static_cast<int>(s >> 8) % 2001 - 1000) / 500.f;
  }
  return v;
}

static float max_diff(const std::vector<float>& a, const std::vector<float>& b) {
  float d = 0.f;
  for (size_t i = 0; i < a.size(); ++i) d = std::fmax(d, std::fabs(a[i] - b[i]));
  return d;
}

static void test_convolution() {
  const size_t sizes[][2] = {{50, 7}, {201, 9}, {400, 300}, {1000, 33}, {4096, 129}, {17, 40}, {1, 1}};
  for (auto& sz : sizes) {
    const size_t xl = sz[0], hl = sz[1];
    auto x = noise(xl, 1), h = noise(hl, 2);
    std::vector<float> want(xl + hl - 1, 0.f), wantc(xl + hl - 1, 0.f);
    for (size_t i = 0; i < xl; ++i)
      for (size_t j = 0; j < hl; ++j) {
        want[i + j] += x[i] * h[j];
        wantc[i + (hl - 1 - j)] += x[i] * h[j];
      }
    for (int alg = 0; alg < 3; ++alg) {
      std::vector<float> got(xl + hl - 1), gotc(xl + hl - 1);
      VsConvolution* c = vs_convolve_initialize_algorithm(xl, hl, alg);
      vs_convolve(c, x.data(), h.data(), got.data());
      vs_convolve_finalize(c);
      c = vs_cross_correlate_initialize_algorithm(xl, hl, alg);
      vs_cross_correlate(c, x.data(), h.data(), gotc.data());
      vs_convolve_finalize(c);
      CHECK(max_diff(got, want) < 2e-3f);
      CHECK(max_diff(gotc, wantc) < 2e-3f);
    }
    VsConvolution* c = vs_convolve_initialize(xl, hl);
    CHECK(vs_convolve_algorithm(c) == vs_convolve_choose(xl, hl));
    vs_convolve_finalize(c);
  }
  CHECK(vs_convolve_choose(201, 10) == VS_CONV_OVERLAP_SAVE);
  CHECK(vs_convolve_choose(351, 200) == VS_CONV_FFT);
  CHECK(vs_convolve_choose(100, 10) == VS_CONV_BRUTE);
}

static void test_wavelets() {
  CHECK(vs_wavelet_validate_order(VS_WAVELET_DAUBECHIES, 8) == 1);
  CHECK(vs_wavelet_validate_order(VS_WAVELET_DAUBECHIES, 76) == 1);
  CHECK(vs_wavelet_validate_order(VS_WAVELET_DAUBECHIES, 7) == 0);
  CHECK(vs_wavelet_validate_order(VS_WAVELET_COIFLET, 12) == 1);
  CHECK(vs_wavelet_validate_order(VS_WAVELET_COIFLET, 8) == 0);
  CHECK(vs_wavelet_validate_order(VS_WAVELET_SYMLET, 20) == 1);
  float lo[8], hi[8];
  CHECK(vs_wavelet_filters(VS_WAVELET_DAUBECHIES, 8, 1, lo, hi) == 0);
  float sum = 0.f, energy = 0.f;
  for (int i = 0; i < 8; ++i) {
    sum += lo[i];
    energy += lo[i] * lo[i];
    CHECK(std::fabs(hi[7 - i] - ((i & 1) ? lo[i] : -lo[i])) < 1e-7f);
  }
  CHECK(std::fabs(sum - std::sqrt(2.f)) < 1e-5f && std::fabs(energy - 1.f) < 1e-5f);
  CHECK(vs_wavelet_filters(VS_WAVELET_SYMLET, 2, 1, lo, hi) == 0 && std::fabs(lo[0] - 0.5f) < 1e-7f);
  CHECK(vs_wavelet_filters(VS_WAVELET_SYMLET, 2, 0, lo, hi) == 0 && std::fabs(lo[0] - std::sqrt(0.5f)) < 1e-6f);

  // periodic extension with an orthonormal bank keeps the energy; every extension against a direct restatement
  const int n = 64, order = 8;
  auto x = noise(n, 5);
  vs_wavelet_filters(VS_WAVELET_DAUBECHIES, order, 1, lo, hi);
  for (int ext = 0; ext < 4; ++ext) {
    std::vector<float> dhi(n / 2), dlo(n / 2), whi(n / 2), wlo(n / 2);
    vs_wavelet_apply(VS_WAVELET_DAUBECHIES, order, ext, x.data(), n, dhi.data(), dlo.data());
    for (int d = 0; d < n / 2; ++d) {
      float rh = 0.f, rl = 0.f;
      for (int j = 0; j < order; ++j) {
        const int idx = 2 * d + j;
        float v;
        if (idx < n) v = x[idx];
        else if (ext == VS_EXT_PERIODIC) v = x[(idx - n) % n];
        else if (ext == VS_EXT_MIRROR) v = x[n - 1 - (idx - n) % n];
        else if (ext == VS_EXT_CONSTANT) v = x[n - 1];
        else v = 0.f;
        rh += hi[j] * v;
        rl += lo[j] * v;
      }
      whi[d] = rh;
      wlo[d] = rl;
    }
    CHECK(max_diff(dhi, whi) < 1e-5f && max_diff(dlo, wlo) < 1e-5f);
    if (ext == VS_EXT_PERIODIC) {
      float e_in = 0.f, e_out = 0.f;
      for (int i = 0; i < n; ++i) e_in += x[i] * x[i];
      for (int i = 0; i < n / 2; ++i) e_out += dhi[i] * dhi[i] + dlo[i] * dlo[i];
      CHECK(std::fabs(e_in - e_out) < 1e-3f * e_in);
    }
    // level 1 of the stationary transform interleaves the decimated one with its one-sample shift
    std::vector<float> shi(n), slo(n);
    vs_stationary_wavelet_apply(VS_WAVELET_DAUBECHIES, order, 1, ext, x.data(), n, shi.data(), slo.data());
    for (int d = 0; d < n / 2; ++d) CHECK(std::fabs(shi[2 * d] - dhi[d]) < 1e-5f && std::fabs(slo[2 * d] - dlo[d]) < 1e-5f);
    vs_stationary_wavelet_apply(VS_WAVELET_DAUBECHIES, order, 3, ext, x.data(), n, shi.data(), slo.data());
    if (ext == VS_EXT_PERIODIC) {
      float rl = 0.f;
      for (int j = 0; j < order; ++j) rl += lo[j] * x[(5 + 4 * j) % n];
      CHECK(std::fabs(slo[5] - rl) < 1e-5f);
    }
  }
  float* src = vs_wavelet_prepare_array(order, x.data(), n);
  float *a, *b, *c, *d;
  vs_wavelet_recycle_source(order, src, n, &a, &b, &c, &d);
  CHECK(a == src && b == src + 16 && c == src + 32 && d == src + 48);
  vs_free(src);
  float* dst = vs_wavelet_allocate_destination(order, n);
  CHECK(dst != nullptr);
  vs_free(dst);
}

static void test_normalize_peaks() {
  std::vector<uint8_t> img(40 * 30);
  for (size_t i = 0; i < img.size(); ++i) img[i] = static_cast<uint8_t>(10 + (i * 37) % 200);
  uint8_t lo, hi;
  vs_minmax2D(img.data(), 40, 33, 30, &lo, &hi);
  uint8_t wlo = 255, whi = 0;
  for (int y = 0; y < 30; ++y)
    for (int x = 0; x < 33; ++x) {
      wlo = std::min(wlo, img[y * 40 + x]);
      whi = std::max(whi, img[y * 40 + x]);
    }
  CHECK(lo == wlo && hi == whi);
  std::vector<float> out(36 * 30, -7.f);
  vs_normalize2D(img.data(), 40, 33, 30, out.data(), 36);
  float omin = 9.f, omax = -9.f;
  for (int y = 0; y < 30; ++y)
    for (int x = 0; x < 33; ++x) {
      omin = std::fmin(omin, out[y * 36 + x]);
      omax = std::fmax(omax, out[y * 36 + x]);
    }
  CHECK(std::fabs(omin + 1.f) < 1e-6f && std::fabs(omax - 1.f) < 1e-6f);
  CHECK(out[33] == -7.f);   // padding columns untouched
  std::vector<uint8_t> flat(16, 9);
  std::vector<float> fo(16, 3.f);
  vs_normalize2D(flat.data(), 4, 4, 4, fo.data(), 4);
  CHECK(fo[0] == 0.f && fo[15] == 0.f);
  auto x = noise(1000, 3);
  float fl, fh;
  vs_minmax1D(x.data(), 1000, &fl, &fh);
  float wl = x[0], wh = x[0];
  for (float v : x) {
    wl = std::fmin(wl, v);
    wh = std::fmax(wh, v);
  }
  CHECK(fl == wl && fh == wh);

  const float sig[] = {0, 1, 0, -1, 0, 0, 2, 2, 1, 3, -5, 4};
  VsExtremumPoint* pts;
  size_t n;
  vs_detect_peaks(sig, 12, VS_EXTREMUM_BOTH, &pts, &n);
  CHECK(n == 5 && pts[0].position == 1 && pts[1].position == 3 && pts[2].position == 8 && pts[4].value == -5.f);
  vs_free(pts);
  vs_detect_peaks(sig, 12, VS_EXTREMUM_MAXIMUM, &pts, &n);
  CHECK(n == 2 && pts[1].position == 9);
  vs_free(pts);
  const float ramp[] = {0, 1, 2, 3};
  vs_detect_peaks(ramp, 4, VS_EXTREMUM_BOTH, &pts, &n);
  CHECK(n == 0 && pts == nullptr);
}

static void test_basic() {
  void* p = vs_malloc_aligned(100);
  CHECK((reinterpret_cast<uintptr_t>(p) & 63) == 0);
  vs_free(p);
  CHECK(vs_next_highest_power_of_2(100) == 128 && vs_next_highest_power_of_2(128) == 128 &&
        vs_next_highest_power_of_2(1) == 1);
  auto x = noise(100, 7);
  size_t nl;
  float* padded = vs_zeropadding(x.data(), 100, &nl);
  CHECK(nl == 256 && padded[99] == x[99] && padded[100] == 0.f && padded[255] == 0.f);
  CHECK(vs_align_complement_f32(padded) == 0 && vs_align_complement_f32(padded + 1) == 15);
  vs_free(padded);
  int16_t i16[5] = {-32768, -1, 0, 1, 32767};
  float f[5];
  vs_int16_to_float(i16, 5, f);
  CHECK(f[0] == -32768.f && f[4] == 32767.f);
  const float g[4] = {1.9f, -1.9f, 100.5f, -0.2f};
  int16_t o16[4];
  vs_float_to_int16(g, 4, o16);
  CHECK(o16[0] == 1 && o16[1] == -1 && o16[2] == 100 && o16[3] == 0);
  const uint16_t halves[5] = {0x3c00, 0xc000, 0x0001, 0x7bff, 0x0000};
  vs_float16_to_float(halves, 5, f);
  CHECK(f[0] == 1.f && f[1] == -2.f && std::fabs(f[2] - 5.9604645e-8f) < 1e-12f && f[3] == 65504.f && f[4] == 0.f);
  const float ca[4] = {1, 2, 3, 4}, cb[4] = {5, 6, 7, 8};
  float cr[4];
  vs_complex_multiply(ca, cb, 4, cr);
  CHECK(cr[0] == -7.f && cr[1] == 16.f && cr[2] == -11.f && cr[3] == 52.f);
  vs_complex_multiply_conjugate(ca, cb, 4, cr);
  CHECK(cr[0] == 17.f && cr[1] == 4.f);
  const float m1[6] = {1, 2, 3, 4, 5, 6};          // 2 x 3
  const float m2[6] = {7, 8, 9, 10, 11, 12};       // 3 x 2
  float r[4];
  vs_matrix_multiply(m1, m2, 3, 2, 2, 3, r);
  CHECK(r[0] == 58.f && r[1] == 64.f && r[2] == 139.f && r[3] == 154.f);
  const float m2t[6] = {7, 9, 11, 8, 10, 12};      // 2 x 3 = m2^T
  vs_matrix_multiply_transposed(m1, m2t, 3, 2, 3, 2, r);
  CHECK(r[0] == 58.f && r[1] == 64.f && r[2] == 139.f && r[3] == 154.f);
  float z[8] = {1, 0, 0, 0, 0, 0, 0, 0};           // delta -> flat spectrum
  CHECK(vs_fft_complex(z, 4, 1) == 0 && z[0] == 1.f && z[2] == 1.f && z[6] == 1.f && z[1] == 0.f);
  CHECK(vs_fft_complex(z, 3, 1) == -1);
}

// Timing harness (the reference's simd/tests/benchmark.inc compares its SIMD and plain variants; here the three
// convolution algorithms and the transforms are timed against each other): `test_simd --bench`.
template <typename F>
static double time_ms(F&& fn, int repeats) {
  fn();
  const auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < repeats; ++i) fn();
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / repeats;
}

static void bench() {
  const size_t sizes[][2] = {{512, 64}, {4096, 64}, {4096, 1024}, {65536, 512}};
  for (auto& sz : sizes) {
    auto x = noise(sz[0], 1), h = noise(sz[1], 2);
    std::vector<float> out(sz[0] + sz[1] - 1);
    printf("convolve %6zu x %4zu:", sz[0], sz[1]);
    for (int alg = 0; alg < 3; ++alg) {
      VsConvolution* c = vs_convolve_initialize_algorithm(sz[0], sz[1], alg);
      const double ms = time_ms([&] { vs_convolve(c, x.data(), h.data(), out.data()); }, alg == 0 && sz[0] > 10000 ? 1 : 5);
      printf("  %s %8.3f ms", alg == 0 ? "direct" : (alg == 1 ? "fft" : "overlap-save"), ms);
      vs_convolve_finalize(c);
    }
    printf("   (heuristic picks %d)\n", vs_convolve_choose(sz[0], sz[1]));
  }
  auto x = noise(1 << 20, 3);
  std::vector<float> hi(1 << 19), lo(1 << 19), shi(1 << 20), slo(1 << 20);
  for (int order : {8, 20, 76})
    printf("dwt daubechies %2d, 1M samples: %7.3f ms\n", order,
           time_ms([&] { vs_wavelet_apply(VS_WAVELET_DAUBECHIES, order, VS_EXT_PERIODIC, x.data(), x.size(), hi.data(), lo.data()); }, 5));
  printf("swt daubechies 8 level 3, 1M samples: %7.3f ms\n",
         time_ms([&] { vs_stationary_wavelet_apply(VS_WAVELET_DAUBECHIES, 8, 3, VS_EXT_PERIODIC, x.data(), x.size(), shi.data(), slo.data()); }, 3));
  auto a = noise(512 * 512, 4), b = noise(512 * 512, 5);
  std::vector<float> c(512 * 512);
  printf("matrix_multiply 512^3: %7.3f ms, transposed: %7.3f ms\n",
         time_ms([&] { vs_matrix_multiply(a.data(), b.data(), 512, 512, 512, 512, c.data()); }, 3),
         time_ms([&] { vs_matrix_multiply_transposed(a.data(), b.data(), 512, 512, 512, 512, c.data()); }, 3));
  VsExtremumPoint* pts;
  size_t n;
  printf("detect_peaks 1M samples: %7.3f ms\n", time_ms([&] {
           vs_detect_peaks(x.data(), x.size(), VS_EXTREMUM_BOTH, &pts, &n);
           vs_free(pts);
         }, 5));
}

int main(int argc, char** argv) {
  if (argc > 1 && std::string(argv[1]) == "--bench") {
    bench();
    return 0;
  }
  test_basic();
  test_convolution();
  test_wavelets();
  test_normalize_peaks();
  if (g_failed) {
    printf("%d checks failed\n", g_failed);
    return 1;
  }
  printf("all simd checks passed\n");
  return 0;
}
